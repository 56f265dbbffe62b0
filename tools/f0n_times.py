"""per-kernel device times of one F0Ntrain forward (torch.profiler). usage: python tools/f0n_times.py [bf16|fp32]"""
import os, sys, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import styletts2_b200 as st2, synthetic as synth
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
p = st2.F0NPredictor(**synth.F0N_CFG)
p.load_state_dict(synth.synth_f0n_state_dict(synth.F0N_CFG), strict=True)
p = p.cuda().eval(); p.precision = prec
inp = synth.synth_f0n_inputs(synth.F0N_CFG, 32, 400)
x, s = inp["x"].cuda(), inp["s"].cuda()
for _ in range(3): p.F0Ntrain(x, s)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    p.F0Ntrain(x, s); torch.cuda.synchronize()
agg = collections.OrderedDict()
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        a = agg.setdefault(e.name[:70], [0, 0.0]); a[0] += 1; a[1] += e.device_time
tot = sum(a[1] for a in agg.values())
print(f"F0Ntrain B=32 x 400 frames {prec}: {sum(a[0] for a in agg.values())} device activities, sum {tot/1e3:.3f} ms")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]): print(f"  {k:70s} n={a[0]:3d} {a[1]/1e3:8.3f} ms {100*a[1]/tot:5.1f}%")

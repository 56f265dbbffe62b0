"""per-kernel device times of one warm Decoder.forward (torch.profiler / CUPTI, no ncu).
usage: python tools/kernel_times.py [c1|c2|c4] [bf16|fp32] [--list]"""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import styletts2_b200 as st2
import synthetic as synth
from bench import WORKLOADS

args = [a for a in sys.argv[1:] if not a.startswith("--")]
wl = args[0] if len(args) > 0 else "c2"
prec = args[1] if len(args) > 1 else "bf16"
kind, B, T, desc = WORKLOADS[wl]
cfg = synth.ISTFTNET_CFG if kind == "istftnet" else synth.HIFIGAN_CFG
dec = (st2.IstftnetDecoder if kind == "istftnet" else st2.HifiganDecoder)(**{k: v for k, v in cfg.items() if k != "kind"})
dec.load_state_dict(synth.synth_state_dict(cfg, seed=0)); dec = dec.cuda().eval(); dec.precision = prec
inp = {k: v.cuda() for k, v in synth.synth_inputs(cfg, B, T, seed=1234, with_noise=False).items()}
for _ in range(3):
    out = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"])
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    out = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"])
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
agg = collections.OrderedDict()
tot = 0.0
for e in evs:
    n = e.name.split("(")[0][:60]
    d = e.device_time if hasattr(e, "device_time") else e.cuda_time
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += d; tot += d
print(f"{desc}: {len(evs)} device activities, sum {tot / 1e3:.3f} ms")
for n, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {n:60s} n={c:4d} {d / 1e3:9.3f} ms {100 * d / tot:5.1f}%")
if "--list" in sys.argv:
    for i, e in enumerate(evs):
        d = e.device_time if hasattr(e, "device_time") else e.cuda_time
        if d > 20:
            print(i, e.name.split("(")[0][:50], round(d, 1))

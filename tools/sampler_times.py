"""per-kernel device times of one warm style-sampler run (torch.profiler / CUPTI, no ncu).
usage: python tools/sampler_times.py [B] [n_tokens] [steps] [bf16|fp32] [scale]"""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import styletts2_b200 as st2
import synthetic as synth

a = sys.argv[1:]
B, n, steps = int(a[0]) if a else 64, int(a[1]) if len(a) > 1 else 128, int(a[2]) if len(a) > 2 else 5
prec, scale = a[3] if len(a) > 3 else "bf16", float(a[4]) if len(a) > 4 else 1.0
cfg = synth.SAMPLER_CFG
m = st2.StyleSampler(**cfg)
m.load_state_dict(synth.synth_sampler_state_dict(cfg, seed=0)); m = m.cuda().eval(); m.precision = prec
inp = {k: v.cuda() for k, v in synth.synth_sampler_inputs(cfg, B, n, steps).items()}
run = lambda: m(inp["noise"], num_steps=steps, embedding=inp["embedding"], features=inp["features"], embedding_scale=scale)
for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); run(); e1.record(); torch.cuda.synchronize()
print(f"B={B} n={n} steps={steps} {prec} scale={scale}: {e0.elapsed_time(e1):.3f} ms wall (events), {m.last_launches} launches")
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    run()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
agg, tot = collections.OrderedDict(), 0.0
for e in evs:
    nme = e.name.split("(")[0][:60]
    d = e.device_time if hasattr(e, "device_time") else e.cuda_time
    x = agg.setdefault(nme, [0, 0.0]); x[0] += 1; x[1] += d; tot += d
print(f"{len(evs)} device activities, sum {tot / 1e3:.3f} ms")
for nme, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {nme:60s} n={c:4d} {d / 1e3:9.3f} ms {100 * d / tot:5.1f}%  avg {d / c:8.1f} us")
if os.environ.get("LIST"):
    evs.sort(key=lambda e: e.time_range.start)
    for i, e in enumerate(evs[: int(os.environ["LIST"])]):
        d = e.device_time if hasattr(e, "device_time") else e.cuda_time
        print(i, e.name.split("(")[0][:50], round(d, 1))

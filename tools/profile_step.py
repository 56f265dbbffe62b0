"""one profiled Decoder.forward (after 2 warm-ups) bracketed by cudaProfilerStart/Stop, for
`ncu --profile-from-start off`.  usage: python tools/profile_step.py [c1|c2|c4] [bf16|fp32]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import styletts2_b200 as st2
import synthetic as synth
from bench import WORKLOADS

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16"
kind, B, T, desc = WORKLOADS[wl]
cfg = synth.ISTFTNET_CFG if kind == "istftnet" else synth.HIFIGAN_CFG
dec = (st2.IstftnetDecoder if kind == "istftnet" else st2.HifiganDecoder)(**{k: v for k, v in cfg.items() if k != "kind"})
dec.load_state_dict(synth.synth_state_dict(cfg, seed=0)); dec = dec.cuda().eval(); dec.precision = prec
inp = {k: v.cuda() for k, v in synth.synth_inputs(cfg, B, T, seed=1234, with_noise=False).items()}
for _ in range(2):
    out = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"])
torch.cuda.synchronize()
torch.cuda.profiler.start()
out = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok", desc, tuple(out.shape), "launches", dec.last_launches, "finite", bool(torch.isfinite(out).all()))

"""per-stage SNR of the CUDA paths against the fp32 oracle (localises precision loss)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import styletts2_b200 as st2
import synthetic as synth
from oracle import decoder_oracle as orc

kind = sys.argv[1] if len(sys.argv) > 1 else "hifigan"
B, T = 2, int(sys.argv[2]) if len(sys.argv) > 2 else 120
cfg = synth.ISTFTNET_CFG if kind == "istftnet" else synth.HIFIGAN_CFG
sd = synth.synth_state_dict(cfg, seed=0)
dec = (st2.IstftnetDecoder if kind == "istftnet" else st2.HifiganDecoder)(**{k: v for k, v in cfg.items() if k != "kind"})
dec.load_state_dict(sd); dec = dec.cuda().eval()
inp = synth.synth_inputs(cfg, B, T, seed=4321)
hs = orc.source_module(inp["F0_curve"], inp["noise"], orc.Params(sd), 300)
kw = {}
if kind == "istftnet":
    mag, ph = orc.stft_transform(hs, 20, 5); kw["har"] = torch.cat([mag, ph], 1)
else:
    kw["har_source"] = hs
taps = {}
ref = orc.decoder_forward(sd, cfg, inp["asr"], inp["F0_curve"], inp["N"], inp["s"], taps=taps, **kw)
args = [inp[k].cuda() for k in ("asr", "F0_curve", "N", "s")]
ckw = {k: v.cuda() for k, v in kw.items()}
for prec in ("fp32", "bf16"):
    w, t = dec(*args, precision=prec, return_taps=True, **ckw)
    line = [f"{prec}: wav SNR {orc.snr_db(ref, w.cpu()):.1f} dB"]
    for name in sorted(t):
        r = taps[name]
        line.append(f"{name} {orc.snr_db(r, t[name].cpu()):.1f} (rms {r.pow(2).mean().sqrt().item():.3g}, absmax {r.abs().max().item():.3g})")
    print(" | ".join(line))

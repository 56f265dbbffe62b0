"""Generate golden vectors from the REAL reference (run in the build container only).

TEST INFRASTRUCTURE ONLY.  Imports ``/root/reference/Modules/{istftnet,hifigan}.py``
unmodified, loads the synthetic state dict of ``synthetic.py`` into the reference
``Decoder``, captures the harmonic excitation with a forward hook on
``generator.m_source`` (no edits to reference files), and writes

  tests/golden/keys_<kind>.json        state-dict names + shapes of the reference
  tests/golden/init_stats_<kind>.json  per-tensor [std, abs-max, mean] of the reference's OWN
                                       ``torch.manual_seed(0); Decoder(...)`` state dict (pins the
                                       distributions of synthetic.synth_state_dict(init="reference"))
  tests/golden/<kind>_T<T>.npz         inputs, injected excitation, wav, a few taps; weights =
                                       synth_state_dict(init="reference") (the reference's real init regime)
  tests/golden/<kind>_legacy_T<T>.npz  the same with init="legacy" (round-1 regime: N(0, 0.01) resblock
                                       weights, perturbed weight_g and Snake alpha)

``/root/reference`` does not exist on the GPU box, which is why the vectors are
committed.  Usage:  python oracle/make_golden.py
"""
from __future__ import annotations

import json
import os
import sys
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.environ.get("ST2_REFERENCE", "/root/reference")

import synthetic as synth  # noqa: E402
from oracle import decoder_oracle as orc  # noqa: E402


def build_reference(cfg):
    sys.path.insert(0, REF)
    warnings.filterwarnings("ignore")
    if cfg["kind"] == "istftnet":
        from Modules.istftnet import Decoder
    else:
        from Modules.hifigan import Decoder
    kw = {k: v for k, v in cfg.items() if k != "kind"}
    torch.manual_seed(0)
    return Decoder(**kw).eval()


def run_reference(dec, inp, cfg):
    """Forward with the SineGen noise draw replaced by inp['noise'] (hook on l_sin_gen is
    not possible without re-running it, so we recompute m_source's output through the
    reference's own SineGen with torch.randn_like patched for that call only)."""
    cap = {}

    def hook(_m, _i, out):
        cap["har_source"] = out[0].detach().clone()
        return out

    h = dec.generator.m_source.register_forward_hook(hook)
    taps = {}
    hooks = [h]

    def tap(name):
        def f(_m, _i, out):
            taps[name] = out.detach().clone()
        return f

    hooks.append(dec.encode.register_forward_hook(tap("encode")))
    for i in range(4):
        hooks.append(dec.decode[i].register_forward_hook(tap(f"decode{i}")))
    if cfg["kind"] == "istftnet":
        hooks.append(dec.generator.conv_post.register_forward_hook(tap("conv_post")))
    # patch randn_like so that the (B,L,9) draw returns our injected noise; the (B,L,1)
    # draw at istftnet.py:295 is unused downstream.
    orig = torch.randn_like
    noise = inp["noise"]

    def fake_randn_like(t, *a, **k):
        if t.dim() == 3 and t.shape[-1] == 9:
            return noise.to(t.dtype)
        return orig(t, *a, **k)

    torch.randn_like = fake_randn_like
    try:
        with torch.no_grad():
            wav = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"])
    finally:
        torch.randn_like = orig
        for hh in hooks:
            hh.remove()
    if cfg["kind"] == "istftnet":
        hs = cap["har_source"].transpose(1, 2).squeeze(1)
        mag, ph = dec.generator.stft.transform(hs)
        taps["har"] = torch.cat([mag, ph], dim=1)
    taps["har_source"] = cap["har_source"][:, :, 0]
    return wav, taps


def main():
    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gold, exist_ok=True)
    report = {}
    for cfg in (synth.ISTFTNET_CFG, synth.HIFIGAN_CFG):
        kind = cfg["kind"]
        dec = build_reference(cfg)
        ref_sd = dec.state_dict()
        spec = synth.param_spec(cfg)
        ref_keys = {k: list(v.shape) for k, v in ref_sd.items()}
        my_keys = {n: list(s) for n, s, *_ in spec}
        assert ref_keys == my_keys, (set(ref_keys) ^ set(my_keys),
                                     [k for k in ref_keys if k in my_keys and ref_keys[k] != my_keys[k]])
        with open(os.path.join(gold, f"keys_{kind}.json"), "w") as f:
            json.dump(ref_keys, f, indent=0, sort_keys=True)
        stats = {k: [float(f"{x:.6g}") for x in v] for k, v in synth.init_stats(ref_sd).items()}
        with open(os.path.join(gold, f"init_stats_{kind}.json"), "w") as f:
            json.dump(stats, f, indent=0, sort_keys=True)
        # the oracle on the reference's own seed-0 weights (not committed: 53 M parameters)
        inp = synth.synth_inputs(cfg, B=1, T=40, seed=999)
        own_sd = {k: v.clone() for k, v in ref_sd.items()}
        wav, taps = run_reference(dec, inp, cfg)
        kw = {"har": taps["har"]} if "har" in taps else {"har_source": taps["har_source"]}
        o = orc.decoder_forward(own_sd, cfg, inp["asr"], inp["F0_curve"], inp["N"], inp["s"], **kw)
        report[f"{kind}_own_seed0_init_T40"] = dict(err=float((o - wav).abs().max()), wav_absmax=float(wav.abs().max()),
                                                     wav_rms=float(wav.pow(2).mean().sqrt()))
        print(kind, "own init", report[f"{kind}_own_seed0_init_T40"], flush=True)
        for init in ("reference", "legacy"):
            sd = synth.synth_state_dict(cfg, seed=0, init=init)
            dec.load_state_dict(sd, strict=True)
            tag = kind if init == "reference" else kind + "_legacy"
            for T, full in ((8, True), (40, False)):
                inp = synth.synth_inputs(cfg, B=2, T=T, seed=1234 + T)
                wav, taps = run_reference(dec, inp, cfg)
                out = {k: v.numpy() for k, v in inp.items()}
                out["wav"] = wav.numpy()
                out["har_source"] = taps["har_source"].numpy()
                if full:
                    for k, v in taps.items():
                        out["tap_" + k] = v.numpy()
                elif "har" in taps:
                    out["tap_har"] = taps["har"].numpy()
                np.savez_compressed(os.path.join(gold, f"{tag}_T{T}.npz"), **out)
                # oracle vs reference, three excitation entry points
                o_noise = orc.decoder_forward(sd, cfg, inp["asr"], inp["F0_curve"], inp["N"], inp["s"],
                                              noise=inp["noise"])
                o_hs = orc.decoder_forward(sd, cfg, inp["asr"], inp["F0_curve"], inp["N"], inp["s"],
                                           har_source=taps["har_source"])
                rep = dict(noise=float((o_noise - wav).abs().max()),
                           har_source=float((o_hs - wav).abs().max()),
                           wav_absmax=float(wav.abs().max()), wav_rms=float(wav.pow(2).mean().sqrt()))
                if "har" in taps:
                    o_har = orc.decoder_forward(sd, cfg, inp["asr"], inp["F0_curve"], inp["N"], inp["s"],
                                                har=taps["har"])
                    rep["har"] = float((o_har - wav).abs().max())
                report[f"{tag}_T{T}"] = rep
                print(tag, T, rep, flush=True)
    with open(os.path.join(gold, "oracle_vs_reference.json"), "w") as f:
        json.dump(report, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()

"""Golden vectors of the style sampler from the REAL reference (build container only).

TEST INFRASTRUCTURE ONLY.  Imports /root/reference/Modules/diffusion/{modules,sampler}.py unmodified
(einops_exts, absent here, is replaced by a three-line ``rearrange_many``), builds the denoiser and the
sampler exactly as models.py:1057-1089 and Demo/Inference_LibriTTS.ipynb cell 14 do, loads the synthetic
state dict (synthetic.py) and records inputs / outputs under tests/golden/sampler_*.npz; the per-step
``randn_like`` draws are captured by wrapping torch.randn_like.   Usage: python oracle/make_golden_sampler.py
"""
from __future__ import annotations

import json
import os
import sys
import types
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.environ.get("ST2_REFERENCE", "/root/reference")

import synthetic as synth  # noqa: E402
from oracle import sampler_oracle as sorc  # noqa: E402


def reference_modules():
    sys.path.insert(0, REF)
    warnings.filterwarnings("ignore")
    if "einops_exts" not in sys.modules:
        from einops import rearrange
        m = types.ModuleType("einops_exts")
        m.rearrange_many = lambda ts, pattern, **kw: tuple(rearrange(t, pattern, **kw) for t in ts)
        sys.modules["einops_exts"] = m
    from Modules.diffusion import modules as mod, sampler as smp  # noqa: E402
    return mod, smp


def build(cfg, sd, mod, smp):
    kw = dict(num_layers=cfg["num_layers"], num_heads=cfg["num_heads"], head_features=cfg["head_features"],
              multiplier=cfg["multiplier"])
    if cfg["context_features"]:
        net = mod.StyleTransformer1d(channels=cfg["channels"], context_embedding_features=cfg["context_embedding_features"],
                                     context_features=cfg["context_features"], **kw)
    else:
        net = mod.Transformer1d(channels=cfg["channels"], context_embedding_features=cfg["context_embedding_features"], **kw)
    ref_keys = {k: list(v.shape) for k, v in net.state_dict().items()}
    assert ref_keys == {k: list(v.shape) for k, v in sd.items()}, "synthetic spec differs from the reference key set"
    net.load_state_dict(sd, strict=True)
    net.eval()
    diff = smp.KDiffusion(net=net, sigma_distribution=smp.LogNormalDistribution(mean=-3.0, std=1.0), sigma_data=0.2,
                          dynamic_threshold=0.0)
    sampler = smp.DiffusionSampler(diff, sampler=smp.ADPM2Sampler(),
                                   sigma_schedule=smp.KarrasSchedule(sigma_min=0.0001, sigma_max=3.0, rho=9.0), clamp=False)
    return net, diff, sampler, ref_keys


def main():
    mod, smp = reference_modules()
    gold = os.path.join(ROOT, "tests", "golden")
    report = {}
    cases = [("style", synth.SAMPLER_CFG, 2, 11, 5, 1.0), ("style_cfg", synth.SAMPLER_CFG, 2, 37, 4, 2.0),
             ("lj", synth.SAMPLER_CFG_LJ, 1, 23, 5, 1.0)]
    for name, cfg, B, n, steps, scale in cases:
        sd = synth.synth_sampler_state_dict(cfg, seed=0)
        net, diff, sampler, ref_keys = build(cfg, sd, mod, smp)
        json.dump(ref_keys, open(os.path.join(gold, f"keys_sampler_{'style' if cfg['context_features'] else 'lj'}.json"), "w"),
                  indent=0, sort_keys=True)
        inp = synth.synth_sampler_inputs(cfg, B, n, steps, seed=11 + n)
        draws = iter(inp["step_noise"])
        orig = torch.randn_like
        torch.randn_like = lambda x, **kw: next(draws).to(x.dtype)
        try:
            with torch.no_grad():
                kw = dict(embedding=inp["embedding"], embedding_scale=scale, num_steps=steps)
                if cfg["context_features"]:
                    kw["features"] = inp["features"]
                out = sampler(noise=inp["noise"], **kw)
                # one denoiser evaluation (KDiffusion.denoise_fn) for the network-level parity test
                den = diff.denoise_fn(inp["noise"] * 0.7, sigma=0.45, embedding=inp["embedding"], embedding_scale=scale,
                                      **({"features": inp["features"]} if cfg["context_features"] else {}))
        finally:
            torch.randn_like = orig
        mine = sorc.sample(sd, cfg, inp["noise"], inp["embedding"], inp["features"], steps, scale, inp["step_noise"])
        den_o = sorc.denoise(sorc.Net(sd, cfg), inp["noise"] * 0.7, 0.45, 0.2, embedding=inp["embedding"],
                             features=inp["features"], embedding_scale=scale)
        report[name] = {"oracle_vs_reference_max_abs": float((mine - out).abs().max()),
                        "denoise_oracle_vs_reference_max_abs": float((den_o - den).abs().max()),
                        "out_absmax": float(out.abs().max())}
        arrs = dict(noise=inp["noise"].numpy(), embedding=inp["embedding"].numpy(), step_noise=inp["step_noise"].numpy(),
                    out=out.numpy(), denoised=den.numpy(), num_steps=np.int32(steps), embedding_scale=np.float32(scale))
        if cfg["context_features"]:
            arrs["features"] = inp["features"].numpy()
        np.savez_compressed(os.path.join(gold, f"sampler_{name}.npz"), **arrs)
    print(json.dumps(report, indent=1))
    json.dump(report, open(os.path.join(gold, "oracle_vs_reference_sampler.json"), "w"), indent=1)


if __name__ == "__main__":
    main()

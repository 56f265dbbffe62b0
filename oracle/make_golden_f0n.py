"""Golden vectors of ``ProsodyPredictor.F0Ntrain`` from the REAL reference (build container only).

TEST INFRASTRUCTURE ONLY.  Imports /root/reference/models.py unmodified; the two packages it imports
that are absent here (munch, einops_exts) are not used by ProsodyPredictor and are stubbed with empty
modules.  The synthetic state dict (synthetic.py) is loaded into the reference module (strict on the
F0Ntrain keys) and inputs/outputs are written to tests/golden/f0n_T<T>.npz, the key set to
tests/golden/keys_f0n.json.   Usage: python oracle/make_golden_f0n.py
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
from oracle import predictor_oracle as porc  # noqa: E402


class _Stub(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return type(name, (), {})


def reference_predictor(cfg):
    sys.path.insert(0, REF)
    warnings.filterwarnings("ignore")
    for name in ("munch", "einops_exts"):
        if name not in sys.modules:
            m = _Stub(name)
            m.__path__ = []
            sys.modules[name] = m
    from models import ProsodyPredictor  # noqa: E402
    torch.manual_seed(0)
    return ProsodyPredictor(style_dim=cfg["style_dim"], d_hid=cfg["d_hid"], nlayers=3, max_dur=50, dropout=0.2).eval()


def main():
    cfg = synth.F0N_CFG
    ref = reference_predictor(cfg)
    sd = synth.synth_f0n_state_dict(cfg, seed=0)
    full = ref.state_dict()
    mine = {k: list(v.shape) for k, v in full.items() if k.split(".")[0] in ("shared", "F0", "N", "F0_proj", "N_proj")}
    assert mine == {k: list(v.shape) for k, v in sd.items()}, "synthetic spec differs from the reference key set"
    full.update(sd)
    ref.load_state_dict(full, strict=True)
    gold = os.path.join(ROOT, "tests", "golden")
    json.dump(mine, open(os.path.join(gold, "keys_f0n.json"), "w"), indent=0, sort_keys=True)
    report = {}
    for B, T in ((2, 20), (3, 51)):
        inp = synth.synth_f0n_inputs(cfg, B, T, seed=77 + T)
        with torch.no_grad():
            F0, N = ref.F0Ntrain(inp["x"], inp["s"])
            lstm, _ = ref.shared(inp["x"].transpose(-1, -2))
        o0, o1 = porc.f0n_forward(sd, inp["x"], inp["s"])
        report[f"T{T}"] = {"oracle_vs_reference_max_abs": float(max((o0 - F0).abs().max(), (o1 - N).abs().max())),
                           "F0_absmax": float(F0.abs().max())}
        np.savez_compressed(os.path.join(gold, f"f0n_T{T}.npz"), x=inp["x"].numpy(), s=inp["s"].numpy(),
                            F0=F0.numpy(), N=N.numpy(), lstm=lstm.numpy())
    print(json.dumps(report, indent=1))
    json.dump(report, open(os.path.join(gold, "oracle_vs_reference_f0n.json"), "w"), indent=1)


if __name__ == "__main__":
    main()

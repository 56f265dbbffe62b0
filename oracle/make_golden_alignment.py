"""Pins oracle/alignment_oracle.py on the reference itself (TEST INFRASTRUCTURE ONLY; run in the build container).

The duration-alignment step has no function in the reference: it is a block of statements inside ``inference()`` of
/root/reference/Demo/Inference_LibriTTS.ipynb (code cell 16; the LJSpeech notebook's cell 17 holds the same loop without
the HiFi-GAN shift).  This script reads the notebook JSON, cuts the statements from ``pred_aln_trg = torch.zeros(`` to the
last ``asr = asr_new`` out of that cell VERBATIM, and exec()s them on seeded durations / encoder outputs with the few
names they use bound to stand-ins (``model.predictor.F0Ntrain`` returns None: it is another row's subject).  Inputs and
the block's outputs (``pred_aln_trg``, ``en``, ``asr``) go to tests/golden/alignment_notebook.npz; tests compare both the
oracle restatement and the CUDA kernels with them.

usage: python oracle/make_golden_alignment.py
"""
import json
import os
import textwrap
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NOTEBOOK = "/root/reference/Demo/Inference_LibriTTS.ipynb"
FIRST, LAST = "pred_aln_trg = torch.zeros(", "asr = asr_new"


def notebook_block() -> str:
    nb = json.load(open(NOTEBOOK))
    cells = [c for c in nb["cells"] if c["cell_type"] == "code" and any(l.startswith("def inference(") for l in c["source"])]
    assert len(cells) == 1, "expected exactly one cell defining inference()"
    lines = "".join(cells[0]["source"]).split("\n")
    lo = next(i for i, l in enumerate(lines) if l.strip().startswith(FIRST))
    hi = max(i for i, l in enumerate(lines) if l.strip() == LAST)
    return textwrap.dedent("\n".join(lines[lo:hi + 1]))


def run_block(block: str, dur: torch.Tensor, d: torch.Tensor, t_en: torch.Tensor, decoder_type: str):
    """dur [Ltext] (float, as torch.round leaves it), d [1, Ltext, Cd], t_en [1, Ct, Ltext]"""
    ns = {
        "torch": torch, "device": "cpu",
        "input_lengths": torch.LongTensor([dur.shape[0]]), "pred_dur": dur.clone(), "d": d, "t_en": t_en, "s": None,
        "model_params": types.SimpleNamespace(decoder=types.SimpleNamespace(type=decoder_type)),
        "model": types.SimpleNamespace(predictor=types.SimpleNamespace(F0Ntrain=lambda en, s: (None, None))),
    }
    exec(block, ns)
    return ns["pred_aln_trg"], ns["en"], ns["asr"]


def main():
    block = notebook_block()
    out = {}
    cases = [("a", 7, 0), ("b", 33, 1), ("c", 1, 2), ("d", 120, 3)]
    for name, ltext, seed in cases:
        g = torch.Generator().manual_seed(seed)
        dur = torch.randint(1, 9, (ltext,), generator=g).float()   # round(...).clamp(min=1)
        if name == "b":
            dur[-1] = 1                                            # the notebook's "pred_dur[-1] = 1" branch
        d = torch.randn(1, ltext, 24, generator=g)
        t_en = torch.randn(1, 16, ltext, generator=g)
        out[f"{name}_dur"] = dur.numpy()
        out[f"{name}_d"] = d.numpy()
        out[f"{name}_t_en"] = t_en.numpy()
        for typ in ("istftnet", "hifigan"):
            aln, en, asr = run_block(block, dur, d, t_en, typ)
            out[f"{name}_{typ}_aln"] = aln.numpy()
            out[f"{name}_{typ}_en"] = en.numpy()
            out[f"{name}_{typ}_asr"] = asr.numpy()
    path = os.path.join(ROOT, "tests", "golden", "alignment_notebook.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(block.split("\n")), "notebook lines executed")
    print(block)


if __name__ == "__main__":
    main()

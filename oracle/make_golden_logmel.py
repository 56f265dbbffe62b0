"""Pins oracle.decoder_oracle.log_mel (and through it the CUDA st2_log_mel kernel) on the reference's own mel front end
(TEST INFRASTRUCTURE ONLY; run in the build container, needs torchaudio).

/root/reference/meldataset.py cannot be imported here (librosa / soundfile are absent), so the statements that define the
front end -- ``to_mel = torchaudio.transforms.MelSpectrogram(...)``, ``mean, std = -4, 4`` and ``def preprocess(wave)``
(meldataset.py:35-43) -- are cut out of the file VERBATIM and exec()'d with torch / torchaudio in scope.  A seeded wave
goes through ``preprocess`` and the result is stored in tests/golden/logmel.npz.

usage: python oracle/make_golden_logmel.py
"""
import os

import numpy as np
import torch
import torchaudio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/meldataset.py"


def reference_preprocess():
    lines = open(SRC).read().split("\n")
    lo = next(i for i, l in enumerate(lines) if l.startswith("to_mel = torchaudio.transforms.MelSpectrogram("))
    hi = next(i for i, l in enumerate(lines) if l.startswith("def preprocess(wave)"))
    end = next(i for i in range(hi + 1, len(lines)) if lines[i].strip().startswith("return "))
    ns = {"torch": torch, "torchaudio": torchaudio}
    exec("\n".join(lines[lo:end + 1]), ns)
    return ns["preprocess"]


def main():
    preprocess = reference_preprocess()
    out = {}
    for name, n, seed in (("a", 6000, 0), ("b", 24000, 1), ("c", 1500, 2)):
        g = torch.Generator().manual_seed(seed)
        t = torch.arange(n) / 24000.0
        wave = (0.3 * torch.sin(2 * torch.pi * 220.0 * t) + 0.1 * torch.sin(2 * torch.pi * 3100.0 * t + 1.0)
                + 0.05 * torch.randn(n, generator=g)).float()
        mel = preprocess(wave.numpy())          # [1, 80, n // 300 + 1]
        out[f"{name}_wave"] = wave.numpy()
        out[f"{name}_mel"] = mel.numpy()
    path = os.path.join(ROOT, "tests", "golden", "logmel.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()

"""Duration-based alignment on the device.

The reference expands token-rate features to frame rate with a one-hot matrix built in a Python loop
on the host and a matmul (``Demo/Inference_LJSpeech.ipynb`` cell 17, ``Inference_LibriTTS.ipynb``
cell 16: ``pred_aln_trg[i, c:c+dur[i]] = 1``, ``asr = t_en @ pred_aln_trg``, and for HiFi-GAN a
one-frame right shift).  Here the durations become a per-frame token index on the device
(:func:`alignment_indices`); :func:`expand` gathers with it, and ``Decoder.forward(t_en, F0, N, s,
alignment=idx)`` fuses the gather into the decoder's input load so ``asr`` is never materialised.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ._cabi import lib, check, ptr
import ctypes as C


def alignment_indices(dur: torch.Tensor, frames: Optional[int] = None, shift: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    """``dur`` [B, Ltext] (or [Ltext]) integer durations on a CUDA device -> (``idx`` int32 [B, frames],
    ``totals`` int32 [B]).  ``frames`` defaults to the largest total (one host sync to read it); frames
    past an item's own total hold -1 (zero columns of the one-hot matrix).  ``shift`` = the HiFi-GAN
    one-frame right shift of the LibriTTS notebook."""
    if not dur.is_cuda:
        raise RuntimeError("alignment_indices needs a CUDA tensor; there is no CPU path")
    d = dur.reshape(1, -1) if dur.dim() == 1 else dur
    d = d.to(torch.int32).contiguous()
    B, L = d.shape
    if frames is None:
        frames = int(d.clamp(min=0).sum(dim=1).max().item())
    frames = max(int(frames), 1)
    idx = torch.empty(B, frames, dtype=torch.int32, device=d.device)
    totals = torch.empty(B, dtype=torch.int32, device=d.device)
    with torch.cuda.device(d.device):
        check(lib.st2_alignment_indices(ptr(d), B, L, frames, 1 if shift else 0, ptr(idx), ptr(totals),
                                        C.c_void_p(torch.cuda.current_stream(d.device).cuda_stream)),
              "st2_alignment_indices")
    return idx, totals


def expand(x: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """``x @ one_hot`` as a gather: ``x`` [B, C, Ltext] fp32, ``idx`` int32 [B, frames] -> [B, C, frames]."""
    if not x.is_cuda or not idx.is_cuda:
        raise RuntimeError("expand needs CUDA tensors; there is no CPU path")
    if x.dim() != 3 or idx.dim() != 2 or idx.shape[0] != x.shape[0]:
        raise RuntimeError("expected x [B, C, Ltext] and idx [B, frames]")
    x_ = x.to(torch.float32).contiguous()
    idx_ = idx.to(torch.int32).contiguous()
    B, Cn, L = x_.shape
    out = torch.empty(B, Cn, idx_.shape[1], dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(lib.st2_expand_alignment(ptr(x_), ptr(idx_), B, Cn, L, idx_.shape[1], ptr(out),
                                       C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)),
              "st2_expand_alignment")
    return out

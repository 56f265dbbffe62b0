"""Output stage after the decoder.

The reference ends with ``out.squeeze().cpu().numpy()[..., :-50]`` (``Demo/Inference_LibriTTS.ipynb``
cell 16): a full fp32 download followed by a host-side trim.  :func:`to_pcm16` trims and converts to
16-bit PCM on the device, which halves the bytes that cross PCIe; ``HostPipeline`` can use it for
its downloads.  The conversion is ``clip(rint(x * 32767), -32768, 32767)`` (round half to even).

:func:`log_mel` is the reference's mel front end (``meldataset.py:35-43``) on the device: the metric the bf16 path's
quality is reported in (log-mel L1) can be computed without a download."""
from __future__ import annotations

import ctypes as C

import torch

from ._cabi import lib, check, ptr


def to_pcm16(wav: torch.Tensor, trim_tail: int = 0) -> torch.Tensor:
    """``wav`` [B, 1, L] or [B, L] fp32 on a CUDA device -> int16 [B, L - trim_tail]."""
    if not wav.is_cuda:
        raise RuntimeError("to_pcm16 needs a CUDA tensor; there is no CPU path")
    w = wav.reshape(wav.shape[0], -1).to(torch.float32).contiguous()
    B, L = w.shape
    if not 0 <= trim_tail < L:
        raise ValueError("trim_tail must be in [0, L)")
    out = torch.empty(B, L - trim_tail, dtype=torch.int16, device=w.device)
    with torch.cuda.device(w.device):
        check(lib.st2_wave_to_pcm16(ptr(w), B, L, trim_tail, ptr(out),
                                    C.c_void_p(torch.cuda.current_stream(w.device).cuda_stream)), "st2_wave_to_pcm16")
    return out


def log_mel(wav: torch.Tensor) -> torch.Tensor:
    """``wav`` [B, 1, L] or [B, L] fp32 on a CUDA device -> normalised log-mel [B, 80, L // 300 + 1] exactly as
    ``meldataset.preprocess`` computes it (torchaudio MelSpectrogram(n_mels=80, n_fft=2048, win_length=1200,
    hop_length=300), ``(log(1e-5 + mel) + 4) / 4``)."""
    if not wav.is_cuda:
        raise RuntimeError("log_mel needs a CUDA tensor; there is no CPU path")
    w = wav.reshape(wav.shape[0], -1).to(torch.float32).contiguous()
    B, L = w.shape
    if L <= 1024:
        raise ValueError("log_mel needs more than 1024 samples per item (reflect padding of n_fft / 2)")
    out = torch.empty(B, 80, L // 300 + 1, dtype=torch.float32, device=w.device)
    with torch.cuda.device(w.device):
        check(lib.st2_log_mel(ptr(w), B, L, ptr(out), C.c_void_p(torch.cuda.current_stream(w.device).cuda_stream)),
              "st2_log_mel")
    return out

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


def wav_frame(pcm16, sample_rate: int = 24000) -> bytes:
    """RIFF / WAVE framing of one utterance (row N4: "fp32 -> int16 PCM / WAV framing"): ``pcm16`` = int16 samples, mono,
    as a 1-D torch tensor (any device: it is downloaded here) or anything ``memoryview`` accepts -> the bytes of a
    .wav file (44-byte PCM header + little-endian samples), what ``IPython.display.Audio`` / ``soundfile.write`` of
    the notebooks produce from the float waveform.  24 kHz is the reference's rate (``Demo/Inference_LibriTTS.ipynb``)."""
    import struct
    import sys
    if isinstance(pcm16, torch.Tensor):
        if pcm16.dtype != torch.int16 or pcm16.dim() != 1:
            raise ValueError("wav_frame takes one utterance of int16 samples (to_pcm16(...)[i])")
        data = pcm16.detach().cpu().contiguous().numpy().tobytes()
    else:
        mv = memoryview(pcm16)
        if mv.itemsize != 2:
            raise ValueError("wav_frame takes 16-bit samples")
        data = mv.tobytes()
    if sys.byteorder != "little":  # WAVE is little-endian
        import array
        a = array.array("h")
        a.frombytes(data)
        a.byteswap()
        data = a.tobytes()
    if not 0 < sample_rate < 2 ** 31 or len(data) > 2 ** 32 - 45:
        raise ValueError("wav_frame: sample rate / length out of range for a RIFF file")
    n_ch, bits = 1, 16
    hdr = struct.pack("<4sI4s4sIHHIIHH4sI", b"RIFF", 36 + len(data), b"WAVE", b"fmt ", 16, 1, n_ch, sample_rate,
                      sample_rate * n_ch * bits // 8, n_ch * bits // 8, bits, b"data", len(data))
    return hdr + data

"""``ProsodyPredictor.F0Ntrain`` on the B200: the step right before the
decoder, producing the ``F0_curve`` and ``N`` it consumes.

Mirrors the reference boundary (/root/reference/models.py:550-590, 638-654): the module holds the
``shared`` bidirectional LSTM, the ``F0`` / ``N`` stacks of three ``AdainResBlk1d`` and the two 1x1
projections under the reference's state-dict names, and exposes ``F0Ntrain(x, s)``.  The rest of
``ProsodyPredictor`` (duration encoder, duration LSTM) is outside the hot path and not built; load a
full ``predictor`` state dict with :meth:`load_predictor_state_dict`, which picks the F0Ntrain keys.
Forward-only, CUDA sm_100 tensors only, no CPU fallback (same rules as the decoder modules).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import torch

from . import _cabi
from ._cabi import lib, check, ptr
from .decoder import _DecoderBase

_F0N_PREFIXES = ("shared.", "F0.", "N.", "F0_proj.", "N_proj.")


class F0NPredictor(_DecoderBase):
    """``predictor.F0Ntrain(en, s) -> (F0 [B, 2T], N [B, 2T])`` (models.py:651-654)."""
    KIND = _cabi.KIND_F0N

    def __init__(self, style_dim: int = 128, d_hid: int = 512):
        super().__init__()
        cfg = _cabi.St2Config()
        cfg.kind = _cabi.KIND_F0N
        cfg.dim_in, cfg.style_dim = int(d_hid), int(style_dim)
        self.d_hid, self.style_dim = int(d_hid), int(style_dim)
        self._setup(cfg)

    @torch.no_grad()
    def _reset_parameters(self):
        super()._reset_parameters()
        k = 1.0 / math.sqrt(self.d_hid // 2)  # nn.LSTM: every parameter U(-1/sqrt(hidden), 1/sqrt(hidden))
        for name, p in self.named_parameters():
            if name.startswith("shared."):
                p.uniform_(-k, k)

    def load_predictor_state_dict(self, sd, strict: bool = True):
        """load the F0Ntrain subset of a full ``ProsodyPredictor`` state dict (models.py:550-590)"""
        sub = {}
        for k, v in sd.items():
            kk = k[7:] if k.startswith("module.") else k
            if kk.startswith(_F0N_PREFIXES):
                sub[kk] = v
        return self.load_state_dict(sub, strict=strict)

    def forward(self, x: torch.Tensor, s: torch.Tensor, precision: Optional[str] = None):
        return self.F0Ntrain(x, s, precision=precision)

    def F0Ntrain(self, x: torch.Tensor, s: torch.Tensor, precision: Optional[str] = None):
        if self.training:
            raise RuntimeError("styletts2_b200.F0NPredictor is forward-only: call .eval() first")
        if not x.is_cuda:
            raise RuntimeError("styletts2_b200.F0NPredictor needs CUDA tensors (sm_100a); there is no CPU path")
        if x.dim() != 3 or s.dim() != 2 or x.shape[0] != s.shape[0]:
            raise RuntimeError("expected x [B, d_hid + style, T] and s [B, style]")
        B, Cin, T = x.shape
        if Cin != self.d_hid + self.style_dim or s.shape[1] != self.style_dim:
            raise RuntimeError(f"input_size mismatch: x has {Cin} channels, the LSTM expects "
                               f"{self.d_hid + self.style_dim}; s has {s.shape[1]}, expected {self.style_dim}")
        prec = precision or self.precision
        prec_id = {"fp32": _cabi.PREC_FP32, "bf16": _cabi.PREC_BF16, "fp32tc": _cabi.PREC_FP32_TC}[prec]
        if prec == "fp32tc" and not self.fp32_tc:
            self.fp32_tc = True   # packs the 3 x bf16 blobs at the next weight upload
        dev = x.device
        x_ = x.to(dev, torch.float32).contiguous()
        s_ = s.to(dev, torch.float32).contiguous()
        with self._lock, torch.cuda.device(dev):
            h = self._handle_for(dev)
            need = lib.st2_f0n_workspace_bytes(h, B, T, prec_id)
            if need < 0:
                check(int(need), "st2_f0n_workspace_bytes")
            ws = torch.empty(int(need), dtype=torch.uint8, device=dev)
            f0 = torch.empty(B, 2 * T, dtype=torch.float32, device=dev)
            n = torch.empty(B, 2 * T, dtype=torch.float32, device=dev)
            stream = torch.cuda.current_stream(dev).cuda_stream
            check(lib.st2_f0n_forward(h, ptr(x_), ptr(s_), ptr(f0), ptr(n), ptr(ws), int(need), B, T, prec_id,
                                      C.c_void_p(stream)), "st2_f0n_forward")
            self.last_launches = int(lib.st2_decoder_last_launches(h))
        return f0, n

"""CPU restatement of ``ProsodyPredictor.F0Ntrain`` (/root/reference/models.py:638-654).

TEST INFRASTRUCTURE ONLY (imported by tests/, smoke() and bench.py's cpu_baseline leg; never by the
product path).  Pinned against the real reference by oracle/make_golden_f0n.py ->
tests/golden/f0n_T*.npz (tests/test_oracle_golden.py).
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.nn.functional as F

from .decoder_oracle import Params, adain_resblk1d


def lstm_direction(x: torch.Tensor, w_ih, w_hh, b_ih, b_hh, reverse: bool) -> torch.Tensor:
    """one direction of nn.LSTM (batch_first, zero initial state), gate order i, f, g, o
    (models.py:562-564 builds it; torch.nn.LSTM semantics).  x [B, T, In] -> [B, T, H]."""
    B, T, _ = x.shape
    H = w_hh.shape[1]
    pre = F.linear(x, w_ih, b_ih + b_hh)  # [B, T, 4H]
    h = x.new_zeros(B, H)
    c = x.new_zeros(B, H)
    out = x.new_zeros(B, T, H)
    steps = range(T - 1, -1, -1) if reverse else range(T)
    for t in steps:
        g = pre[:, t] + F.linear(h, w_hh)
        i, f, gg, o = g.chunk(4, dim=1)
        c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
        h = torch.sigmoid(o) * torch.tanh(c)
        out[:, t] = h
    return out


def shared_lstm_aten(x_ncl: torch.Tensor, P: Params) -> torch.Tensor:
    """The same layer the way the reference runs it: ATen's fused ``nn.LSTM(d_hid + style, d_hid // 2, 1,
    batch_first=True, bidirectional=True)`` (models.py:562-564, 639) with the state dict's weights.  Used by the timed
    CPU baseline of ``bench.py --workload f0n`` (the Python time loop below understates the reference's CPU speed) and
    checked against the explicit recurrence in tests/test_oracle_golden.py."""
    w_ih = P.get("shared.weight_ih_l0")
    lstm = torch.nn.LSTM(w_ih.shape[1], w_ih.shape[0] // 4, 1, batch_first=True, bidirectional=True).to(w_ih.dtype)
    with torch.no_grad():
        for name, prm in lstm.named_parameters():
            prm.copy_(P.get("shared." + name))
        return lstm(x_ncl.transpose(1, 2))[0]


def shared_lstm(x_ncl: torch.Tensor, P: Params, aten: bool = False) -> torch.Tensor:
    """``x, _ = self.shared(x.transpose(-1, -2))`` (models.py:639): [B, In, T] -> [B, T, 2H]."""
    if aten:
        return shared_lstm_aten(x_ncl, P)
    x = x_ncl.transpose(1, 2)
    outs = []
    for sfx, rev in (("", False), ("_reverse", True)):
        outs.append(lstm_direction(x, P.get("shared.weight_ih_l0" + sfx), P.get("shared.weight_hh_l0" + sfx),
                                   P.get("shared.bias_ih_l0" + sfx), P.get("shared.bias_hh_l0" + sfx), rev))
    return torch.cat(outs, dim=2)


def f0n_forward(sd: Dict[str, torch.Tensor], x: torch.Tensor, s: torch.Tensor,
                dtype=torch.float32, aten_lstm: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    """``F0Ntrain(x, s)`` (models.py:638-654): x [B, d_hid + style, T], s [B, style] -> F0, N [B, 2T].
    ``aten_lstm``: run the recurrence through ``torch.nn.LSTM`` like the reference (timed baseline)."""
    P = Params(sd, dtype)
    x, s = x.to(dtype), s.to(dtype)
    h = shared_lstm(x, P, aten_lstm).transpose(1, 2)  # [B, 2H, T]
    outs = []
    for br in ("F0", "N"):
        y = h
        for i, up in enumerate((False, True, False)):
            y = adain_resblk1d(y, s, P, f"{br}.{i}", up)
        y = F.conv1d(y, P.get(br + "_proj.weight"), P.get(br + "_proj.bias"))
        outs.append(y.squeeze(1))
    return outs[0], outs[1]

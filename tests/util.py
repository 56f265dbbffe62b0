"""Shared helpers for the GPU parity tests: run one fused conv through the C ABI and restate
the same op with torch on the CPU in float64 (conv_params.h semantics)."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch
import torch.nn.functional as F

from styletts2_b200 import _cabi
from styletts2_b200._cabi import lib, check


def conv_reference(x, ss, act, slope, alpha, w_kcn, bias, k, dil, pad, stride, Lout, res, res_shift,
                   out_scale, y_old):
    """x [B, Lin, cin] (values as stored), w_kcn [k, cin, n]; everything float64 on CPU."""
    x = x.double()
    if ss is not None:
        x = x * ss[:, None, :, 0].double() + ss[:, None, :, 1].double()
    if act == _cabi.ACT_LRELU:
        x = torch.where(x > 0, x, x * slope)
    elif act == _cabi.ACT_SNAKE:
        a = alpha.double()[None, None, :]
        x = x + (1 / a) * torch.sin(a * x) ** 2
    w = w_kcn.double().permute(2, 1, 0).contiguous()  # [n, cin, k]
    y = F.conv1d(x.transpose(1, 2), w, None if bias is None else bias.double(), stride=stride,
                 padding=pad, dilation=dil).transpose(1, 2)  # [B, L, n]
    if y.shape[1] < Lout:
        y = F.pad(y, (0, 0, 0, Lout - y.shape[1]))
    y = y[:, :Lout]
    if res is not None:
        idx = torch.arange(Lout) >> res_shift
        y = y + res.double()[:, idx]
    y = y * out_scale
    if y_old is not None:
        y = y + y_old.double()
    return y


def run_conv(impl, x, w_kcn, bias, *, k, dil=1, pad=0, stride=1, Lout=None, ss=None, act=_cabi.ACT_NONE,
             slope=0.0, alpha=None, res=None, res_shift=0, out_scale=1.0, y_old=None, want_stats=False,
             out_dtype=None, cin_pad=None):
    """x: cuda tensor [B, Lin, ld] (float32 or bfloat16), contraction over the first cin columns.
    Returns (y [B, Lout, n] in out dtype, stats [B, n, 2] or None)."""
    dev = x.device
    B, Lin, ld = x.shape
    kk, cin, n = w_kcn.shape
    assert kk == k
    umma = impl == "umma"
    out_dtype = out_dtype or x.dtype
    if Lout is None:
        Lout = (Lin + 2 * pad - dil * (k - 1) - 1) // stride + 1
    p = _cabi.St2ConvParams()
    p.B = B
    p.x = x.data_ptr(); p.x_dtype = _cabi.DT_BF16 if x.dtype == torch.bfloat16 else _cabi.DT_F32
    p.x_bstride = Lin * ld; p.ldx = ld; p.Lin = Lin
    keep = []
    if umma:
        cin_pad = cin_pad or (cin + 63) // 64 * 64
        assert ld >= cin_pad
        elems = lib.st2_umma_packed_elems(k, cin_pad, n)
        packed = torch.empty(elems, dtype=torch.int16)
        wk = w_kcn.float().contiguous().cpu()
        check(lib.st2_umma_pack_weights(wk.data_ptr(), k, cin, cin_pad, n, packed.data_ptr()))
        wd = packed.to(dev)
        p.cin = cin_pad
    else:
        wd = w_kcn.float().contiguous().to(dev)
        p.cin = cin
    keep.append(wd)
    p.w = wd.data_ptr()
    if ss is not None:
        ssd = ss.float().contiguous().to(dev)  # [B, C, 2]
        keep.append(ssd)
        p.ss = ssd.data_ptr(); p.ss_stride = ssd.shape[1]
    p.act = act; p.slope = slope
    if alpha is not None:
        ad = alpha.float().contiguous().to(dev)
        if umma and ad.numel() < p.cin:
            ad = torch.cat([ad, torch.ones(p.cin - ad.numel(), device=dev)])
        keep.append(ad)
        p.alpha = ad.data_ptr()
    p.k, p.dil, p.pad, p.stride = k, dil, pad, stride
    if bias is not None:
        bd = bias.float().contiguous().to(dev); keep.append(bd); p.bias = bd.data_ptr()
    p.n_cols = n
    if y_old is not None:
        y = y_old.to(dev, out_dtype).contiguous().clone()
    else:
        y = torch.full((B, Lout, n), float("nan"), device=dev, dtype=out_dtype)
    p.y = y.data_ptr(); p.y_dtype = _cabi.DT_BF16 if out_dtype == torch.bfloat16 else _cabi.DT_F32
    p.y_bstride = Lout * n; p.ldy = n; p.Lout = Lout
    if res is not None:
        rd = res.to(dev, out_dtype).contiguous(); keep.append(rd)
        p.res = rd.data_ptr(); p.res_bstride = rd.shape[1] * rd.shape[2]; p.ldres = rd.shape[2]
        p.res_shift = res_shift
    p.out_scale = out_scale
    p.accumulate = 1 if y_old is not None else 0
    stats = None
    if want_stats:
        tr = lib.st2_conv_tile_rows(1 if umma else 0)
        nt = (Lout + tr - 1) // tr
        stats = torch.full((B, nt, n, 2), float("nan"), device=dev, dtype=torch.float32)
        p.stats = stats.data_ptr(); p.n_tiles = nt
    fn = lib.st2_conv_umma if umma else lib.st2_conv_simt
    check(fn(C.byref(p), None), f"st2_conv_{impl}")
    torch.cuda.synchronize()
    return y, (None if stats is None else stats.sum(dim=1))


def snr_db(ref, test):
    ref, test = ref.double().flatten(), test.double().flatten()
    return float(10 * torch.log10((ref ** 2).sum() / ((ref - test) ** 2).sum().clamp_min(1e-300)))


def wrap_phase_err(a, b):
    d = (a - b).double()
    return (torch.atan2(torch.sin(d), torch.cos(d))).abs()

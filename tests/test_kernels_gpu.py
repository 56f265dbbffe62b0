"""GPU parity tests of the single kernels, called through the C ABI (include/st2_decoder.h).

fp32 FFMA conv: against a float64 torch restatement of conv_params.h (tolerance 2e-5 relative).
tcgen05 conv : against the same restatement fed with bf16-rounded operands (what the tensor core
               multiplies), tolerance = one bf16 ulp of the output + fp32 accumulation noise.
source/STFT/iSTFT: against the oracle (oracle/decoder_oracle.py), which is pinned to the reference.
"""
import ctypes as C
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from styletts2_b200 import _cabi  # noqa: E402
from styletts2_b200._cabi import lib, check  # noqa: E402
from tests.util import conv_reference, run_conv, snr_db, wrap_phase_err  # noqa: E402


def _mk_case(B, Lin, cin, n, k, seed, ld=None, with_ss=True, act=_cabi.ACT_SNAKE):
    g = torch.Generator().manual_seed(seed)
    ld = ld or cin
    x = torch.randn(B, Lin, ld, generator=g)
    w = torch.randn(k, cin, n, generator=g) / math.sqrt(cin * k)
    bias = torch.randn(n, generator=g) * 0.1
    ss = None
    if with_ss:
        ss = torch.stack([1 + 0.5 * (torch.rand(B, ld, generator=g) - 0.5),
                          0.5 * (torch.rand(B, ld, generator=g) - 0.5)], dim=-1)
    alpha = 1 + 0.4 * (torch.rand(ld, generator=g) - 0.5) if act == _cabi.ACT_SNAKE else None
    return x, w, bias, ss, alpha


SIMT_CASES = [
    # B, Lin, cin, n, k, dil, stride, act
    (2, 100, 16, 64, 3, 1, 1, _cabi.ACT_SNAKE),
    (2, 150, 22, 40, 12, 1, 6, _cabi.ACT_NONE),      # noise_convs[0]: 22 -> C, k12 s6 p3
    (1, 301, 128, 22, 7, 1, 1, _cabi.ACT_LRELU),     # conv_post
    (2, 77, 130, 100, 11, 5, 1, _cabi.ACT_SNAKE),    # ragged everything, widest halo
    (1, 600, 1, 16, 60, 1, 30, _cabi.ACT_NONE),      # hifigan noise_convs[0]: 1 -> C, k60 s30
    (2, 64, 514, 96, 3, 1, 1, _cabi.ACT_LRELU),      # encode.conv1 width
]


@pytest.mark.parametrize("case", SIMT_CASES)
def test_conv_simt_fp32(case):
    B, Lin, cin, n, k, dil, stride, act = case
    x, w, bias, ss, alpha = _mk_case(B, Lin, cin, n, k, seed=hash(case) & 0xffff, act=act)
    pad = (stride + 1) // 2 if stride > 1 else dil * (k - 1) // 2
    Lout = (Lin + 2 * pad - dil * (k - 1) - 1) // stride + 1
    g = torch.Generator().manual_seed(7)
    res = torch.randn(B, Lout, n, generator=g)
    ref = conv_reference(x, ss, act, 0.2, alpha, w, bias, k, dil, pad, stride, Lout, res, 0, 0.75, None)
    y, st = run_conv("simt", x.cuda(), w, bias, k=k, dil=dil, pad=pad, stride=stride, ss=ss, act=act,
                     slope=0.2, alpha=alpha, res=res, out_scale=0.75, want_stats=True)
    err = (y.cpu().double() - ref).abs().max().item()
    assert err < 2e-5 * max(1.0, ref.abs().max().item()), err
    s_ref = torch.stack([ref.sum(1), (ref ** 2).sum(1)], dim=-1)
    assert torch.allclose(st.cpu().double(), s_ref, rtol=1e-4, atol=1e-3)


def test_conv_simt_accumulate_and_upsampled_residual():
    B, Lin, cin, n, k = 2, 90, 32, 48, 3
    x, w, bias, ss, alpha = _mk_case(B, Lin, cin, n, k, seed=5, act=_cabi.ACT_LRELU)
    g = torch.Generator().manual_seed(9)
    res = torch.randn(B, Lin // 2, n, generator=g)
    y_old = torch.randn(B, Lin, n, generator=g)
    ref = conv_reference(x, ss, _cabi.ACT_LRELU, 0.2, None, w, bias, k, 1, 1, 1, Lin, res, 1,
                         1 / math.sqrt(2), y_old)
    y, _ = run_conv("simt", x.cuda(), w, bias, k=k, pad=1, ss=ss, act=_cabi.ACT_LRELU, slope=0.2, res=res,
                    res_shift=1, out_scale=1 / math.sqrt(2), y_old=y_old)
    assert (y.cpu().double() - ref).abs().max().item() < 2e-5


def _bf16_round(t):
    return t.to(torch.bfloat16).to(torch.float32)


def _umma_reference(xb, ss, act, slope, alpha, w, bias, k, dil, pad, Lout, res, res_shift, out_scale, y_old):
    """what the tensor-core path computes: bf16(act(affine(x_bf16))) x bf16(w), fp32-ish accumulate."""
    x = xb.float()
    if ss is not None:
        x = x * ss[:, None, :, 0] + ss[:, None, :, 1]
    if act == _cabi.ACT_LRELU:
        x = torch.where(x > 0, x, x * slope)
    elif act == _cabi.ACT_SNAKE:
        a = alpha[None, None, :]
        x = x + (1 / a) * torch.sin(a * x) ** 2
    x = _bf16_round(x)
    return conv_reference(x, None, _cabi.ACT_NONE, 0.0, None, _bf16_round(w), bias, k, dil, pad, 1, Lout,
                          res, res_shift, out_scale, y_old)


UMMA_CASES = [
    # B, Lin, cin, n, k, dil, act, with_res
    (1, 256, 64, 128, 1, 1, _cabi.ACT_NONE, False),    # single tile, single stage, plain GEMM
    (1, 256, 128, 128, 3, 1, _cabi.ACT_NONE, False),   # taps = row shifts of one smem tile
    (2, 700, 128, 128, 7, 3, _cabi.ACT_SNAKE, True),   # stage-2 resblock conv, ragged length
    (2, 1030, 128, 128, 11, 5, _cabi.ACT_SNAKE, True), # widest halo (50 rows)
    (1, 513, 256, 256, 3, 5, _cabi.ACT_SNAKE, True),   # stage-1 width, two N tiles
    (2, 130, 576, 1024, 3, 1, _cabi.ACT_LRELU, False), # encode.conv1: 9 chunks, 8 N tiles
    (3, 40, 1152, 512, 1, 1, _cabi.ACT_NONE, False),   # conv1x1 shortcut
    (1, 1200, 512, 2560, 3, 1, _cabi.ACT_LRELU, True), # ups[0] in polyphase form (N = 10*256)
]


@pytest.mark.parametrize("case", UMMA_CASES)
def test_conv_umma_bf16(case):
    B, Lin, cin, n, k, dil, act, with_res = case
    x, w, bias, ss, alpha = _mk_case(B, Lin, cin, n, k, seed=hash(case) & 0xffff, act=act,
                                     with_ss=act != _cabi.ACT_NONE)
    pad = dil * (k - 1) // 2
    xb = x.to(torch.bfloat16)
    g = torch.Generator().manual_seed(11)
    res = torch.randn(B, Lin, n, generator=g).to(torch.bfloat16) if with_res else None
    ref = _umma_reference(xb, ss, act, 0.2, alpha, w, bias, k, dil, pad, Lin, res, 0, 0.5, None)
    y, st = run_conv("umma", xb.cuda(), w, bias, k=k, dil=dil, pad=pad, ss=ss, act=act, slope=0.2,
                     alpha=alpha, res=res, out_scale=0.5, want_stats=True)
    yd = y.cpu().double()
    assert torch.isfinite(yd).all()
    tol = 2.0 ** -7 * ref.abs() + 3e-3  # one bf16 ulp of the result + rounding flips of the operands
    bad = ((yd - ref).abs() > tol).sum().item()
    assert bad == 0, (bad, (yd - ref).abs().max().item(), snr_db(ref, yd))
    assert snr_db(ref, yd) > 45.0
    s_ref = torch.stack([yd.sum(1), (yd ** 2).sum(1)], dim=-1)  # statistics are of the stored values
    assert torch.allclose(st.cpu().double(), s_ref, rtol=2e-3, atol=2e-2)


@pytest.mark.parametrize("case", [
    # B, Lin, cin, n, k, dil, act, out fp32
    (2, 700, 128, 32, 7, 1, _cabi.ACT_LRELU, True),    # conv_post: 128 -> 22 (padded 32), fp32 rows of 128 B
    (1, 1300, 64, 64, 3, 3, _cabi.ACT_SNAKE, False),   # hifigan stage-3 width: one 64-column tile
    (2, 515, 64, 64, 11, 5, _cabi.ACT_SNAKE, False),   # 64-column tile, widest halo
    (1, 300, 128, 64, 1, 1, _cabi.ACT_NONE, True),     # fp32 out, 64 columns
    (2, 1000, 64, 192, 3, 1, _cabi.ACT_NONE, False),   # three 64-column tiles
])
def test_conv_umma_narrow_tiles(case):
    B, Lin, cin, n, k, dil, act, f32out = case
    x, w, bias, ss, alpha = _mk_case(B, Lin, cin, n, k, seed=hash(case) & 0xffff, act=act,
                                     with_ss=act != _cabi.ACT_NONE)
    pad = dil * (k - 1) // 2
    xb = x.to(torch.bfloat16)
    g = torch.Generator().manual_seed(13)
    res = None if f32out else torch.randn(B, Lin, n, generator=g).to(torch.bfloat16)
    ref = _umma_reference(xb, ss, act, 0.01, alpha, w, bias, k, dil, pad, Lin, res, 0, 1.0, None)
    y, st = run_conv("umma", xb.cuda(), w, bias, k=k, dil=dil, pad=pad, ss=ss, act=act, slope=0.01,
                     alpha=alpha, res=res, out_scale=1.0, want_stats=not f32out,
                     out_dtype=torch.float32 if f32out else None)
    yd = y.cpu().double()
    assert torch.isfinite(yd).all()
    tol = (2.0 ** -7 * ref.abs() if not f32out else 0) + 3e-3
    assert ((yd - ref).abs() > tol).sum().item() == 0, (yd - ref).abs().max().item()
    if not f32out:
        s_ref = torch.stack([yd.sum(1), (yd ** 2).sum(1)], dim=-1)
        assert torch.allclose(st.cpu().double(), s_ref, rtol=2e-3, atol=2e-2)


def test_conv_umma_accumulate_matches_simt():
    B, Lin, cin, n, k, dil = 2, 900, 128, 128, 7, 1
    x, w, bias, ss, alpha = _mk_case(B, Lin, cin, n, k, seed=21)
    xb = x.to(torch.bfloat16).cuda()
    g = torch.Generator().manual_seed(3)
    res = torch.randn(B, Lin, n, generator=g).to(torch.bfloat16)
    y_old = torch.randn(B, Lin, n, generator=g).to(torch.bfloat16)
    kw = dict(k=k, dil=dil, pad=3, ss=ss, act=_cabi.ACT_SNAKE, alpha=alpha, res=res, out_scale=1 / 3,
              y_old=y_old)
    yu, _ = run_conv("umma", xb, w, bias, **kw)
    ys, _ = run_conv("simt", xb, w, bias, **kw)
    assert snr_db(ys.float(), yu.float()) > 40.0


# ------------------------------------------------------------------ source / STFT / iSTFT
def _oracle():
    from oracle import decoder_oracle as orc
    import synthetic as synth
    return orc, synth


@pytest.mark.parametrize("T", [8, 40, 400, 800, 2400])
def test_har_source_matches_oracle(T):
    orc, synth = _oracle()
    cfg = synth.ISTFTNET_CFG
    inp = synth.synth_inputs(cfg, 2, T, seed=77 + T)
    sd = {"generator.m_source.l_linear.weight": torch.tensor([[0.3, -0.2, 0.25, 0.1, -0.15, 0.2, 0.05, -0.1, 0.12]]),
          "generator.m_source.l_linear.bias": torch.tensor([0.03])}
    P = orc.Params(sd)
    ref = orc.source_module(inp["F0_curve"], inp["noise"], P, 300)
    f0 = inp["F0_curve"].cuda().contiguous(); nz = inp["noise"].cuda().contiguous()
    lw = sd["generator.m_source.l_linear.weight"].reshape(-1).cuda().contiguous()
    ph = torch.empty(2 * 9 * T, device="cuda"); out = torch.empty(2, T * 300, device="cuda")
    check(lib.st2_source_har(f0.data_ptr(), nz.data_ptr(), 0, lw.data_ptr(), 0.03, 2, T, 300, ph.data_ptr(),
                             out.data_ptr(), None))
    torch.cuda.synchronize()
    d = (out.cpu() - ref).abs()
    # the phase reaches ~8e4 rad at T=400 (ulp 8e-3) and ~5e5 rad at T=2400 (config C5; ulp 0.03, where the
    # reference's own sine is 8.3e-2 away from an fp64 evaluation): identical fp32 rounding points are required
    assert d.max().item() < 2e-5, (d.max().item(), (d > 1e-5).float().mean().item())


def test_har_source_philox_is_normal_and_seeded():
    T = 40
    f0 = torch.zeros(2, T, device="cuda")  # unvoiced: output = tanh(sum w_h * (0.1/3) * z_h + b)
    lw = torch.zeros(9, device="cuda"); lw[0] = 1.0
    ph = torch.empty(2 * 9 * T, device="cuda")
    outs = []
    for seed in (1, 1, 2):
        out = torch.empty(2, T * 300, device="cuda")
        check(lib.st2_source_har(f0.data_ptr(), None, seed, lw.data_ptr(), 0.0, 2, T, 300, ph.data_ptr(),
                                 out.data_ptr(), None))
        outs.append(out.clone())
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1]) and not torch.equal(outs[0], outs[2])
    z = torch.atanh(outs[0]) / (0.1 / 3)
    assert abs(z.mean().item()) < 0.02 and abs(z.std().item() - 1) < 0.02
    assert (outs[0][0] - outs[0][1]).abs().max().item() > 0  # items draw different noise


@pytest.mark.parametrize("T", [8, 40])
def test_stft_matches_oracle(T):
    orc, synth = _oracle()
    g = torch.Generator().manual_seed(T)
    hs = torch.tanh(torch.randn(2, T * 300, generator=g) * 0.3)
    mag, ph = orc.stft_transform(hs, 20, 5)
    L2 = T * 60 + 1
    out = torch.empty(2, L2, 32, device="cuda")
    check(lib.st2_stft_har(hs.cuda().contiguous().data_ptr(), 2, T * 300, out.data_ptr(), None))
    torch.cuda.synchronize()
    o = out.cpu()
    assert (o[:, :, :11].transpose(1, 2) - mag).abs().max().item() < 2e-6
    # torch.angle is discontinuous at +-pi: compare on the circle, weighted by the magnitude
    perr = wrap_phase_err(o[:, :, 11:22].transpose(1, 2), ph) * mag.double()
    assert perr.max().item() < 5e-6
    assert (o[:, :, 22:] == 0).all()
    # DC / Nyquist: Im == +0 -> angle in {0, +pi}, never -pi (probed reference behaviour)
    assert (o[:, :, 11] >= 0).all() and (o[:, :, 21] >= 0).all()


@pytest.mark.parametrize("T", [8, 40, 800])
def test_stft_hilo_matches_oracle(T):
    """stft_hilo_kernel, the STFT the bf16 (benchmarked) path launches: bf16 (hi, lo) pairs whose sum carries
    ~16 mantissa bits of |X| / angle; padding rows and columns are zero"""
    orc, synth = _oracle()
    g = torch.Generator().manual_seed(100 + T)
    hs = torch.tanh(torch.randn(2, T * 300, generator=g) * 0.3)
    mag, ph = orc.stft_transform(hs, 20, 5)
    L2 = T * 60 + 1
    rows_pad = (L2 + 5) // 6 * 6 + 12
    out = torch.full((2, rows_pad, 64), float("nan"), device="cuda", dtype=torch.bfloat16)
    check(lib.st2_stft_har16(hs.cuda().contiguous().data_ptr(), 2, T * 300, rows_pad, out.data_ptr(), None))
    torch.cuda.synchronize()
    o = out.float().cpu()
    assert (o[:, L2:] == 0).all() and (o[:, :, 22:32] == 0).all() and (o[:, :, 54:] == 0).all()
    v = o[:, :L2, :22] + o[:, :L2, 32:54]
    hi = o[:, :L2, :22]
    assert (hi.to(torch.bfloat16).float() == hi).all()
    m = v[:, :, :11].transpose(1, 2)
    assert ((m - mag).abs() / mag.abs().clamp_min(1e-2)).max().item() < 4e-5     # 2^-16 relative + fp32 DFT noise
    perr = wrap_phase_err(v[:, :, 11:].transpose(1, 2), ph)
    assert (perr * mag.double()).max().item() < 1e-4 * max(1.0, mag.max().item())
    assert (v[:, :, 11] >= 0).all() and (v[:, :, 21] >= 0).all()


@pytest.mark.parametrize("frames", [5, 481, 2401])
def test_istft_head_matches_oracle(frames):
    orc, _ = _oracle()
    g = torch.Generator().manual_seed(frames)
    x = torch.randn(2, 22, frames, generator=g) * 0.7
    ref = orc.istft_inverse(torch.exp(x[:, :11]), torch.sin(x[:, 11:]), 20, 5)
    xcl = torch.zeros(2, frames, 32)
    xcl[:, :, :22] = x.transpose(1, 2)
    wav = torch.empty(2, 5 * (frames - 1), device="cuda")
    check(lib.st2_istft_head(xcl.cuda().contiguous().data_ptr(), 2, frames, wav.data_ptr(), None))
    torch.cuda.synchronize()
    assert (wav.cpu() - ref[:, 0]).abs().max().item() < 5e-6 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("case", [
    # B, Lin, cin, n, k, dil, act, residual, accumulate
    (3, 1000, 128, 128, 7, 3, _cabi.ACT_SNAKE, True, False),    # odd tile count: the last pair computes one tile twice
    (2, 2000, 256, 256, 11, 5, _cabi.ACT_SNAKE, True, False),   # two N tiles, widest halo
    (5, 777, 128, 128, 3, 1, _cabi.ACT_LRELU, False, False),
    (2, 1500, 128, 128, 7, 1, _cabi.ACT_SNAKE, True, True),     # running mean through the identity blocks
    (1, 300, 64, 128, 1, 1, _cabi.ACT_NONE, False, False),      # prologue-free path (TMA bytes straight to the MMAs)
])
def test_conv_umma_cta_pair_mode_is_bit_identical(case, monkeypatch):
    """ST2_UMMA_PAIR=1 runs the same conv with cta_group::2 MMAs over 2-CTA clusters (conv_umma.cu, UmLayout::pair):
    every stored byte and every statistic must equal the single-CTA kernel's."""
    B, Lin, cin, n, k, dil, act, with_res, acc = case
    x, w, bias, ss, alpha = _mk_case(B, Lin, cin, n, k, seed=sum(case[:6]), act=act, with_ss=act != _cabi.ACT_NONE)
    g = torch.Generator().manual_seed(5)
    res = torch.randn(B, Lin, n, generator=g).to(torch.bfloat16) if with_res else None
    y_old = torch.randn(B, Lin, n, generator=g).to(torch.bfloat16) if acc else None
    outs = []
    for mode in ("0", "1"):
        monkeypatch.setenv("ST2_UMMA_PAIR", mode)
        _cabi.check(_cabi.lib.st2_umma_reload_env())       # the switches are cached: read them again
        y, st = run_conv("umma", x.to(torch.bfloat16).cuda(), w, bias, k=k, dil=dil, pad=dil * (k - 1) // 2, ss=ss,
                         act=act, slope=0.1, alpha=alpha, res=res, out_scale=0.5 if acc else 1.0, y_old=y_old,
                         want_stats=not acc)
        torch.cuda.synchronize()
        outs.append((y.float().cpu(), None if st is None else st.cpu()))
    monkeypatch.delenv("ST2_UMMA_PAIR")
    _cabi.check(_cabi.lib.st2_umma_reload_env())
    assert torch.isfinite(outs[0][0]).all()
    assert torch.equal(outs[0][0], outs[1][0])
    if outs[0][1] is not None:
        assert torch.equal(outs[0][1], outs[1][1])


def test_log_mel_kernel_vs_reference_front_end():
    """next row N4: st2_log_mel (2048-point FFT per frame in shared memory + sparse HTK filterbank) against the
    reference's own torchaudio front end (tests/golden/logmel.npz from meldataset.py:35-43) and the oracle; a batch
    of items with ragged lengths goes through one call each"""
    import os
    from styletts2_b200.output import log_mel
    from oracle import decoder_oracle as orc
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "logmel.npz"))
    for n in "abc":
        wave = torch.from_numpy(g[f"{n}_wave"])
        got = log_mel(wave[None].cuda()).cpu()
        ref = torch.from_numpy(g[f"{n}_mel"])
        assert got.shape == ref.shape
        assert (got - ref).abs().max().item() < 2e-4
    w = 0.2 * torch.randn(3, 1, 12000, generator=torch.Generator().manual_seed(4))
    got = log_mel(w.cuda()).cpu().double()
    assert (got - orc.log_mel(w[:, 0])).abs().max().item() < 2e-4
    with pytest.raises(ValueError):
        log_mel(torch.zeros(1, 1000).cuda())


def _mk_job(B, L, cin, n, k, dil, seed, with_res, with_stats, acc=False, act=None):
    """St2ConvParams of one bf16 tcgen05 conv on random data (+ the tensors it points at)"""
    act = _cabi.ACT_SNAKE if act is None else act
    g = torch.Generator().manual_seed(seed)
    dev = "cuda"
    x = torch.randn(B, L, cin, generator=g).to(torch.bfloat16).to(dev)
    w = (torch.randn(k, cin, n, generator=g) / math.sqrt(cin * k)).contiguous()
    packed = torch.empty(lib.st2_umma_packed_elems(k, cin, n), dtype=torch.int16)
    check(lib.st2_umma_pack_weights(w.data_ptr(), k, cin, cin, n, packed.data_ptr()))
    wd = packed.to(dev)
    bias = torch.randn(n, generator=g).to(dev)
    ss = torch.stack([1 + 0.1 * torch.randn(B, cin, generator=g), 0.1 * torch.randn(B, cin, generator=g)], -1).contiguous().to(dev)
    alpha = (0.5 + torch.rand(cin, generator=g)).to(dev)
    y0 = torch.randn(B, L, n, generator=g).to(torch.bfloat16).to(dev)   # y_old for accumulate, overwritten otherwise
    res = torch.randn(B, L, n, generator=g).to(torch.bfloat16).to(dev) if with_res else None
    tr = lib.st2_conv_tile_rows(1)
    nt = (L + tr - 1) // tr
    stats = torch.zeros(B, nt, n, 2, device=dev) if with_stats else None

    def params(y, st):
        p = _cabi.St2ConvParams()
        p.B = B
        p.x = x.data_ptr(); p.x_dtype = _cabi.DT_BF16; p.x_bstride = L * cin; p.ldx = cin; p.Lin = L; p.cin = cin
        p.ss = ss.data_ptr(); p.ss_stride = cin; p.act = act; p.slope = 0.1; p.alpha = alpha.data_ptr()
        p.k, p.dil, p.pad, p.stride = k, dil, dil * (k - 1) // 2, 1
        p.w = wd.data_ptr(); p.bias = bias.data_ptr(); p.n_cols = n
        p.y = y.data_ptr(); p.y_dtype = _cabi.DT_BF16; p.y_bstride = L * n; p.ldy = n; p.Lout = L
        if res is not None:
            p.res = res.data_ptr(); p.res_bstride = L * n; p.ldres = n
        p.out_scale = 0.5 if acc else 1.0
        p.accumulate = 1 if acc else 0
        if st is not None:
            p.stats = st.data_ptr(); p.n_tiles = nt
        return p
    return params, y0, stats, (x, wd, bias, ss, alpha, res)


@pytest.mark.parametrize("case", [
    # B, L, cin = n, [(k, dil, residual, stats, accumulate)]
    (3, 1000, 128, [(3, 1, False, True, False), (11, 1, False, True, False)]),                          # the k3 + k11 pair
    (2, 1500, 128, [(3, 1, True, True, False), (7, 3, True, True, False), (11, 5, True, False, True)]),  # 3 jobs, mixed chunks
    (2, 700, 256, [(3, 1, False, True, False), (11, 3, True, True, False)]),                            # two N tiles
    (1, 300, 128, [(7, 1, True, True, False), (3, 5, False, False, False), (11, 1, False, True, False)]),
])
def test_conv_umma_multi_equals_separate_launches(case):
    """st2_conv_umma_multi: up to 3 same-shaped convs alternate tile by tile inside ONE persistent kernel (conv_umma.cu,
    UmJobs).  Every output byte and every statistic must equal the separately launched convs."""
    B, L, CH, jobs = case
    made = [_mk_job(B, L, CH, CH, k, d, 100 + i, r, s, a) for i, (k, d, r, s, a) in enumerate(jobs)]
    want = []
    for params, y0, stats, keep in made:
        y = y0.clone()
        st = None if stats is None else stats.clone()
        check(lib.st2_conv_umma(C.byref(params(y, st)), None))
        torch.cuda.synchronize()
        want.append((y, st))
    arr = (_cabi.St2ConvParams * len(made))()
    got = []
    for i, (params, y0, stats, keep) in enumerate(made):
        y = y0.clone()
        st = None if stats is None else stats.clone()
        arr[i] = params(y, st)
        got.append((y, st))
    check(lib.st2_conv_umma_multi(arr, len(made), None))
    torch.cuda.synchronize()
    for (yw, sw), (yg, sg) in zip(want, got):
        assert torch.isfinite(yg.float()).all()
        assert torch.equal(yw, yg)
        if sw is not None:
            assert torch.equal(sw, sg)


@pytest.mark.parametrize("case", [
    # B, Lin, n, k, stride: HiFi-GAN's noise_convs on the raw excitation (hifigan.py:301-311), Conv1d(1, C, 2 s, stride s)
    (2, 12000, 256, 60, 30),
    (2, 12000, 128, 12, 6),
    (3, 6000, 64, 4, 2),
    (2, 5000, 32, 1, 1),      # last stage: kernel_size 1
    (1, 130, 64, 4, 2),       # shorter than one block's window
])
def test_conv_thin_one_input_channel(case):
    """conv_simt.cu's thin kernel (cin = 1, bf16 out): a block owns several 64-row statistics tiles and writes their
    sums into the first tile's slot; values against float64, statistics = sums of the stored bf16 values."""
    B, Lin, n, k, stride = case
    g = torch.Generator().manual_seed(Lin + n)
    x = torch.randn(B, Lin, 1, generator=g)
    w = torch.randn(k, 1, n, generator=g) / math.sqrt(k)
    bias = torch.randn(n, generator=g) * 0.1
    pad = (stride + 1) // 2 if stride > 1 else 0
    Lout = (Lin + 2 * pad - (k - 1) - 1) // stride + 1
    ref = conv_reference(x, None, _cabi.ACT_NONE, 0.0, None, w, bias, k, 1, pad, stride, Lout, None, 0, 1.0, None)
    y, st = run_conv("simt", x.cuda(), w, bias, k=k, pad=pad, stride=stride, want_stats=True, out_dtype=torch.bfloat16)
    yd = y.cpu().double()
    assert torch.isfinite(yd).all()
    assert ((yd - ref).abs() <= 2.0 ** -8 * ref.abs() + 1e-6).all()
    s_ref = torch.stack([yd.sum(1), (yd ** 2).sum(1)], dim=-1)
    assert torch.allclose(st.cpu().double(), s_ref, rtol=1e-4, atol=1e-2)

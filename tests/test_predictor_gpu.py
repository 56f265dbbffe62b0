"""GPU parity of ``F0NPredictor.F0Ntrain`` through the module -> C ABI, against the
golden vectors of the REAL reference (tests/golden/f0n_T*.npz) and against the oracle.

Tolerances: fp32 path <= 1e-3 max-abs on both curves (values are O(1) with random-init weights; the
measured error is ~1e-5); bf16 path SNR >= 30 dB against the fp32 oracle."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_CACHE = {}


def _get():
    if "p" not in _CACHE:
        import synthetic as synth
        import styletts2_b200 as st2
        p = st2.F0NPredictor(**synth.F0N_CFG)
        sd = synth.synth_f0n_state_dict(synth.F0N_CFG, seed=0)
        p.load_state_dict(sd, strict=True)
        _CACHE["p"] = (p.cuda().eval(), sd)
    return _CACHE["p"]


@pytest.mark.parametrize("T", [20, 51])
def test_f0n_fp32_vs_reference_golden(T):
    p, sd = _get()
    g = np.load(os.path.join(GOLD, f"f0n_T{T}.npz"))
    g = {k: torch.from_numpy(g[k]) for k in g.files}
    F0, N = p.F0Ntrain(g["x"].cuda(), g["s"].cuda(), precision="fp32")
    assert F0.shape == g["F0"].shape and N.shape == g["N"].shape
    e0, e1 = (F0.cpu() - g["F0"]).abs().max().item(), (N.cpu() - g["N"]).abs().max().item()
    assert max(e0, e1) < 1e-3, (e0, e1)
    assert max(e0, e1) < 1e-4, (e0, e1)  # what the fp32 kernels actually reach


@pytest.mark.parametrize("B,T", [(1, 1), (5, 3), (2, 130), (7, 257)])
def test_f0n_ragged_shapes_vs_oracle(B, T):
    """one-frame input, batches that do not fill an LSTM cluster group (4 items), lengths across conv tiles"""
    from oracle import predictor_oracle as porc
    import synthetic as synth
    p, sd = _get()
    inp = synth.synth_f0n_inputs(synth.F0N_CFG, B, T, seed=900 + B + T)
    r0, r1 = porc.f0n_forward(sd, inp["x"], inp["s"])
    F0, N = p.F0Ntrain(inp["x"].cuda(), inp["s"].cuda(), precision="fp32")
    assert (F0.cpu() - r0).abs().max().item() < 1e-3 and (N.cpu() - r1).abs().max().item() < 1e-3
    b0, b1 = p.F0Ntrain(inp["x"].cuda(), inp["s"].cuda(), precision="bf16")
    assert torch.isfinite(b0).all() and torch.isfinite(b1).all()


def test_f0n_bf16_snr_vs_oracle():
    from oracle import predictor_oracle as porc
    from oracle import decoder_oracle as orc
    import synthetic as synth
    p, sd = _get()
    inp = synth.synth_f0n_inputs(synth.F0N_CFG, 4, 200, seed=31)
    r0, r1 = porc.f0n_forward(sd, inp["x"], inp["s"])
    b0, b1 = p.F0Ntrain(inp["x"].cuda(), inp["s"].cuda(), precision="bf16")
    s0, s1 = orc.snr_db(r0, b0.cpu()), orc.snr_db(r1, b1.cpu())
    print(f"F0Ntrain bf16: SNR F0 {s0:.1f} dB, N {s1:.1f} dB")
    assert min(s0, s1) >= 30.0


def test_f0n_rows_independent_and_deterministic():
    import synthetic as synth
    p, sd = _get()
    inp = synth.synth_f0n_inputs(synth.F0N_CFG, 6, 40, seed=5)
    x, s = inp["x"].cuda(), inp["s"].cuda()
    for prec in ("fp32", "bf16"):
        a0, a1 = p.F0Ntrain(x, s, precision=prec)
        c0, c1 = p.F0Ntrain(x[2:5], s[2:5], precision=prec)
        assert torch.equal(a0[2:5], c0) and torch.equal(a1[2:5], c1), prec
        d0, d1 = p.F0Ntrain(x, s, precision=prec)
        assert torch.equal(a0, d0) and torch.equal(a1, d1), prec


def test_f0n_feeds_the_decoder():
    """the two next-row modules chained the way the reference's inference does (Inference_LJSpeech
    cell 17): F0_pred, N_pred = predictor.F0Ntrain(en, s); decoder(asr, F0_pred, N_pred, ref)"""
    import synthetic as synth
    import styletts2_b200 as st2
    p, sd = _get()
    cfg = synth.ISTFTNET_CFG
    dec = st2.IstftnetDecoder(**{k: v for k, v in cfg.items() if k != "kind"})
    dec.load_state_dict(synth.synth_state_dict(cfg, seed=0), strict=True)
    dec = dec.cuda().eval()
    T = 24
    inp = synth.synth_f0n_inputs(synth.F0N_CFG, 2, T, seed=1)
    F0, N = p.F0Ntrain(inp["x"].cuda(), inp["s"].cuda())
    asr = torch.randn(2, 512, T, device="cuda")
    wav = dec(asr, F0, N, inp["s"].cuda())
    assert wav.shape == (2, 1, 2 * T * 300) and torch.isfinite(wav).all()


def test_f0n_error_behaviour():
    from styletts2_b200._cabi import St2Error
    p, sd = _get()
    with pytest.raises(RuntimeError, match="input_size"):
        p.F0Ntrain(torch.randn(1, 512, 8, device="cuda"), torch.randn(1, 128, device="cuda"))
    with pytest.raises(RuntimeError, match="CUDA"):
        p.F0Ntrain(torch.randn(1, 640, 8), torch.randn(1, 128))
    p.train()
    try:
        with pytest.raises(RuntimeError, match="forward-only"):
            p.F0Ntrain(torch.randn(1, 640, 8, device="cuda"), torch.randn(1, 128, device="cuda"))
    finally:
        p.eval()


@pytest.mark.parametrize("T", [20, 51])
def test_f0n_fp32_tensor_core_path_vs_reference_golden(T):
    """ST2_PREC_FP32_TC for F0Ntrain: the LSTM input projection and the six AdainResBlk1d convs on tcgen05 as 3 x bf16
    products with fp32 accumulation; same golden, <= 1e-3"""
    import styletts2_b200 as st2
    import synthetic as synth
    p = st2.F0NPredictor(**synth.F0N_CFG)
    p.load_state_dict(synth.synth_f0n_state_dict(synth.F0N_CFG, seed=0), strict=True)
    p = p.cuda().eval()
    p.fp32_tc = True
    g = np.load(os.path.join(GOLD, f"f0n_T{T}.npz"))
    F0, N = p.F0Ntrain(torch.from_numpy(g["x"]).cuda(), torch.from_numpy(g["s"]).cuda(), precision="fp32tc")
    e0 = (F0.cpu() - torch.from_numpy(g["F0"])).abs().max().item()
    e1 = (N.cpu() - torch.from_numpy(g["N"])).abs().max().item()
    print(f"F0Ntrain fp32tc T={T}: max-abs F0 {e0:.2e}, N {e1:.2e}")
    assert e0 < 1e-3 and e1 < 1e-3

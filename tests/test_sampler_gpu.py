"""GPU parity of ``StyleSampler`` through the module -> C ABI, against the golden vectors
of the REAL reference (tests/golden/sampler_*.npz, made by oracle/make_golden_sampler.py) and against the oracle.

Tolerances: fp32 path <= 1e-3 max-abs on the sampled style vector and on one denoiser evaluation (values
are O(0.5); measured ~1e-5); bf16 path SNR >= 30 dB against the fp32 oracle (north star's bf16 bar)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_CACHE = {}


def _get(variant="style"):
    if variant not in _CACHE:
        import synthetic as synth
        import styletts2_b200 as st2
        cfg = synth.SAMPLER_CFG if variant == "style" else synth.SAMPLER_CFG_LJ
        m = st2.StyleSampler(**cfg)
        sd = synth.synth_sampler_state_dict(cfg, seed=0)
        m.load_state_dict(sd, strict=True)
        _CACHE[variant] = (m.cuda().eval(), sd, cfg)
    return _CACHE[variant]


def _gold(name):
    g = np.load(os.path.join(GOLD, f"sampler_{name}.npz"))
    return {k: torch.from_numpy(np.asarray(g[k])) for k in g.files}


@pytest.mark.parametrize("name", ["style", "style_cfg", "lj"])
def test_sampler_fp32_vs_reference_golden(name):
    m, sd, cfg = _get("lj" if name == "lj" else "style")
    g = _gold(name)
    feat = g["features"].cuda() if "features" in g else None
    scale, steps = float(g["embedding_scale"]), int(g["num_steps"])
    den = m.denoise_fn((g["noise"] * 0.7).cuda(), 0.45, embedding=g["embedding"].cuda(), embedding_scale=scale,
                       features=feat, precision="fp32")
    e_den = (den.cpu() - g["denoised"]).abs().max().item()
    out = m(g["noise"].cuda(), num_steps=steps, embedding=g["embedding"].cuda(), embedding_scale=scale, features=feat,
            step_noise=g["step_noise"].cuda(), precision="fp32")
    assert out.shape == g["out"].shape
    e_out = (out.cpu() - g["out"]).abs().max().item()
    assert e_den < 1e-3 and e_out < 1e-3, (e_den, e_out)
    assert e_den < 1e-4 and e_out < 2e-4, (e_den, e_out)  # what the fp32 kernels actually reach


@pytest.mark.parametrize("B,n,scale", [(1, 1, 1.0), (3, 7, 1.0), (2, 129, 1.5), (5, 200, 1.0), (1, 512, 1.0)])
def test_sampler_ragged_shapes_vs_oracle(B, n, scale):
    """one token, token counts across the attention key tiles (64) / query blocks (128) / conv tiles (256),
    the maximum length of the fixed embedding, odd batches, with and without guidance"""
    from oracle import sampler_oracle as sorc
    import synthetic as synth
    m, sd, cfg = _get("style")
    inp = synth.synth_sampler_inputs(cfg, B, n, 3, seed=300 + B + n)
    ref = sorc.sample(sd, cfg, inp["noise"], inp["embedding"], inp["features"], 3, scale, inp["step_noise"])
    out = m(inp["noise"].cuda(), num_steps=3, embedding=inp["embedding"].cuda(), embedding_scale=scale,
            features=inp["features"].cuda(), step_noise=inp["step_noise"].cuda(), precision="fp32")
    assert (out.cpu() - ref).abs().max().item() < 1e-3
    b16 = m(inp["noise"].cuda(), num_steps=3, embedding=inp["embedding"].cuda(), embedding_scale=scale,
            features=inp["features"].cuda(), step_noise=inp["step_noise"].cuda(), precision="bf16")
    assert torch.isfinite(b16).all()


@pytest.mark.parametrize("variant,scale", [("style", 1.0), ("style", 2.0), ("lj", 1.0)])
def test_sampler_bf16_snr_vs_oracle(variant, scale):
    from oracle import sampler_oracle as sorc
    from oracle import decoder_oracle as orc
    import synthetic as synth
    m, sd, cfg = _get(variant)
    inp = synth.synth_sampler_inputs(cfg, 4, 96, 5, seed=41)
    feat = inp["features"]
    ref = sorc.sample(sd, cfg, inp["noise"], inp["embedding"], feat, 5, scale, inp["step_noise"])
    out = m(inp["noise"].cuda(), num_steps=5, embedding=inp["embedding"].cuda(), embedding_scale=scale,
            features=None if feat is None else feat.cuda(), step_noise=inp["step_noise"].cuda(), precision="bf16")
    snr = orc.snr_db(ref, out.cpu())
    print(f"sampler bf16 SNR ({variant}, scale {scale}): {snr:.1f} dB")
    assert snr >= 30.0, snr


def test_sampler_rows_independent_and_philox_sharding():
    """items are independent (bitwise), and a shard with item_offset draws the step noise the full batch would"""
    import synthetic as synth
    m, sd, cfg = _get("style")
    inp = synth.synth_sampler_inputs(cfg, 6, 50, 4, seed=9)
    a = {k: (v.cuda() if v is not None else None) for k, v in inp.items()}
    full = m(a["noise"], num_steps=4, embedding=a["embedding"], features=a["features"], precision="bf16", seed=123)
    again = m(a["noise"], num_steps=4, embedding=a["embedding"], features=a["features"], precision="bf16", seed=123)
    assert torch.equal(full, again)
    other = m(a["noise"], num_steps=4, embedding=a["embedding"], features=a["features"], precision="bf16", seed=124)
    assert not torch.equal(full, other)
    shard = m(a["noise"][4:], num_steps=4, embedding=a["embedding"][4:], features=a["features"][4:], precision="bf16",
              seed=123, item_offset=4)
    assert torch.equal(shard, full[4:])


def test_sampler_default_calls_draw_fresh_step_noise():
    """two calls without an explicit seed differ (the reference draws fresh randn_like noise in every
    ADPM2Sampler.step, sampler.py:510); resetting ``.seed`` reproduces the first call"""
    import synthetic as synth
    m, sd, cfg = _get("style")
    inp = synth.synth_sampler_inputs(cfg, 3, 20, 4, seed=2)
    a = {k: (v.cuda() if v is not None else None) for k, v in inp.items()}
    call = lambda: m(a["noise"], num_steps=4, embedding=a["embedding"], features=a["features"], precision="bf16")
    m.seed = 41
    o1, o2 = call(), call()
    assert m.seed == 43 and not torch.equal(o1, o2)
    m.seed = 41
    assert torch.equal(call(), o1)
    # injected step noise or an explicit seed leave the counter alone
    m(a["noise"], num_steps=4, embedding=a["embedding"], features=a["features"], step_noise=a["step_noise"])
    m(a["noise"], num_steps=4, embedding=a["embedding"], features=a["features"], seed=7)
    assert m.seed == 42


def test_sampler_errors_and_prefix_loading():
    import synthetic as synth
    import styletts2_b200 as st2
    m, sd, cfg = _get("style")
    inp = synth.synth_sampler_inputs(cfg, 2, 8, 3)
    with pytest.raises(RuntimeError):  # features missing (modules.py:137)
        m(inp["noise"].cuda(), num_steps=3, embedding=inp["embedding"].cuda())
    with pytest.raises(RuntimeError):  # num_steps missing (sampler.py:577)
        m(inp["noise"].cuda(), embedding=inp["embedding"].cuda(), features=inp["features"].cuda())
    with pytest.raises(RuntimeError):  # CPU tensors
        m(inp["noise"], num_steps=3, embedding=inp["embedding"], features=inp["features"])
    # a full model.diffusion state dict holds the net as diffusion.net.* and unet.* (models.py:1088-1089)
    full = {}
    for k, v in sd.items():
        full["diffusion.net." + k] = v
        full["unet." + k] = v
    m2 = st2.StyleSampler(**cfg)
    m2.load_diffusion_state_dict(full)
    m2 = m2.cuda().eval()
    a = m(inp["noise"].cuda(), num_steps=3, embedding=inp["embedding"].cuda(), features=inp["features"].cuda(),
          step_noise=inp["step_noise"].cuda(), precision="fp32")
    b = m2(inp["noise"].cuda(), num_steps=3, embedding=inp["embedding"].cuda(), features=inp["features"].cuda(),
           step_noise=inp["step_noise"].cuda(), precision="fp32")
    assert torch.equal(a, b)


def test_sampler_feeds_predictor_and_decoder():
    """s_pred[:, 128:] conditions F0Ntrain, s_pred[:, :128] the decoder (Inference_LibriTTS.ipynb cell 16)"""
    import synthetic as synth
    import styletts2_b200 as st2
    m, sd, cfg = _get("style")
    inp = synth.synth_sampler_inputs(cfg, 2, 30, 5, seed=77)
    s_pred = m(inp["noise"].cuda(), num_steps=5, embedding=inp["embedding"].cuda(), features=inp["features"].cuda())
    s_pred = s_pred.squeeze(1)
    ref, s = s_pred[:, :128], s_pred[:, 128:]
    p = st2.F0NPredictor(**synth.F0N_CFG)
    p.load_state_dict(synth.synth_f0n_state_dict(synth.F0N_CFG, seed=0))
    p = p.cuda().eval()
    en = synth.synth_f0n_inputs(synth.F0N_CFG, 2, 20, seed=3)["x"].cuda()
    F0, N = p.F0Ntrain(en, s)
    dcfg = synth.ISTFTNET_CFG
    dec = st2.IstftnetDecoder(**{k: v for k, v in dcfg.items() if k != "kind"})
    dec.load_state_dict(synth.synth_state_dict(dcfg, seed=0))
    dec = dec.cuda().eval()
    asr = torch.randn(2, 512, 20, device="cuda")
    wav = dec(asr, F0, N, ref)
    assert wav.shape == (2, 1, 300 * 40) and torch.isfinite(wav).all()

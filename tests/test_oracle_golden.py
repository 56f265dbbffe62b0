"""CPU: the oracle restatement against the golden vectors the REAL reference produced
(oracle/make_golden.py, run where /root/reference exists).  This is what pins the oracle."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import decoder_oracle as orc
import synthetic as synth
from tests.util_cpu import wrap_phase_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CFGS = {"istftnet": synth.ISTFTNET_CFG, "hifigan": synth.HIFIGAN_CFG}
_SD = {}


INITS = ["reference", "legacy"]   # synthetic.py: the reference's real init regime / the round-1 regime


def sd_of(kind, init="reference"):
    if (kind, init) not in _SD:
        _SD[(kind, init)] = synth.synth_state_dict(CFGS[kind], seed=0, init=init)
    return _SD[(kind, init)]


def gold(kind, T, init="reference"):
    tag = kind if init == "reference" else kind + "_legacy"
    g = np.load(os.path.join(GOLD, f"{tag}_T{T}.npz"))
    return {k: torch.from_numpy(g[k]) for k in g.files}


@pytest.mark.parametrize("kind", ["istftnet", "hifigan"])
def test_param_spec_matches_reference_keys(kind):
    ref = json.load(open(os.path.join(GOLD, f"keys_{kind}.json")))
    mine = {n: list(s) for n, s, *_ in synth.param_spec(CFGS[kind])}
    assert mine == ref
    assert len(ref) == (375 if kind == "istftnet" else 678)


@pytest.mark.parametrize("kind", ["istftnet", "hifigan"])
def test_synthetic_init_matches_the_reference_constructor(kind):
    """synth_state_dict(init="reference") against the per-tensor statistics of the reference's own
    ``torch.manual_seed(0); Decoder(...)`` (tests/golden/init_stats_*.json, oracle/make_golden.py): every
    weight-normed ``weight_v`` is kaiming-uniform (``init_weights``, Modules/utils.py:1-4, writes the derived
    ``.weight`` attribute and so never lands), ``weight_g`` = ||v||, Snake alpha = 1."""
    ref = json.load(open(os.path.join(GOLD, f"init_stats_{kind}.json")))
    mine = synth.init_stats(sd_of(kind))
    assert set(ref) == set(mine)
    n_checked = 0
    for k, (std, amax, mean) in ref.items():
        m_std, m_amax, m_mean = mine[k]
        numel = sd_of(kind)[k].numel()
        if k.endswith("weight_g"):
            assert std < 1e-6 and m_std < 1e-6 and abs(mean - 1) < 1e-6 and abs(m_mean - 1) < 1e-6, k
        elif "alpha" in k:
            assert std == 0 and m_std == 0 and mean == 1 and m_mean == 1, k
        elif numel >= 256:
            tol = 0.05 if numel >= 4096 else 0.2          # both sides are sample estimates
            assert abs(m_std / std - 1) < tol, (k, m_std, std)
            assert abs(m_amax / amax - 1) < tol, (k, m_amax, amax)
            n_checked += 1
        else:
            assert m_amax <= amax * 1.5 + 1e-6 or numel < 16, (k, m_amax, amax)
    assert n_checked > 150
    # the layer the round-1 regime got wrong: kaiming-uniform, not N(0, 0.01)
    assert mine["generator.resblocks.3.convs1.0.weight_v"][0] == pytest.approx(0.0295, abs=0.001)
    legacy = synth.init_stats(sd_of(kind, "legacy"))
    assert legacy["generator.resblocks.3.convs1.0.weight_v"][0] == pytest.approx(0.0100, abs=0.001)


@pytest.mark.parametrize("init", INITS)
@pytest.mark.parametrize("T", [8, 40])
def test_istftnet_oracle_vs_reference_har_injected(T, init):
    g = gold("istftnet", T, init)
    out = orc.decoder_forward(sd_of("istftnet", init), CFGS["istftnet"], g["asr"], g["F0_curve"], g["N"], g["s"],
                              har=g["tap_har"])
    assert out.shape == g["wav"].shape
    assert (out - g["wav"]).abs().max().item() < (5e-6 if init == "reference" else 2e-6)


@pytest.mark.parametrize("init", INITS)
def test_istftnet_oracle_taps_vs_reference(init):
    g = gold("istftnet", 8, init)
    taps = {}
    orc.decoder_forward(sd_of("istftnet", init), CFGS["istftnet"], g["asr"], g["F0_curve"], g["N"], g["s"],
                        noise=g["noise"], taps=taps)
    for name in ("encode", "decode0", "decode1", "decode2", "decode3"):
        assert (taps[name] - g["tap_" + name]).abs().max().item() < 1e-4, name
    assert (taps["har_source"] - g["har_source"]).abs().max().item() < 2e-6
    har, ref = taps["har"], g["tap_har"]
    assert (har[:, :11] - ref[:, :11]).abs().max().item() < 5e-6
    assert (wrap_phase_err(har[:, 11:], ref[:, 11:]) * ref[:, :11].double()).max().item() < 5e-6
    # real bins: angle is 0 or +pi, never -pi
    assert (har[:, 11] >= 0).all() and (har[:, 21] >= 0).all()
    assert (ref[:, 11] >= 0).all() and (ref[:, 21] >= 0).all()


@pytest.mark.parametrize("init", INITS)
@pytest.mark.parametrize("T", [8, 40])
def test_hifigan_oracle_vs_reference(T, init):
    """reference init drives tanh into saturation (rms 0.8, |wav| up to 0.9999): 2e-5 there, 2e-6 legacy"""
    g = gold("hifigan", T, init)
    tol = 5e-5 if init == "reference" else 1e-5
    out = orc.decoder_forward(sd_of("hifigan", init), CFGS["hifigan"], g["asr"], g["F0_curve"], g["N"], g["s"],
                              noise=g["noise"])
    assert (out - g["wav"]).abs().max().item() < tol
    out2 = orc.decoder_forward(sd_of("hifigan", init), CFGS["hifigan"], g["asr"], g["F0_curve"], g["N"], g["s"],
                               har_source=g["har_source"])
    assert (out2 - g["wav"]).abs().max().item() < tol


@pytest.mark.parametrize("kind", ["istftnet", "hifigan"])
def test_library_mode_matches_reference_golden(kind):
    """decoder_forward(library=True) — F.instance_norm, torch.stft, torch.istft: the reference's own torch calls,
    what ``bench.py --impl reference-gpu`` times on the B200 — against the same golden of the real reference"""
    g = gold(kind, 40)
    kw = {"har": g["tap_har"]} if kind == "istftnet" else {"har_source": g["har_source"]}
    out = orc.decoder_forward(sd_of(kind), CFGS[kind], g["asr"], g["F0_curve"], g["N"], g["s"], library=True, **kw)
    assert (out - g["wav"]).abs().max().item() < 5e-5
    hs = g["har_source"]
    m1, p1 = orc.stft_transform(hs, 20, 5, library=True)
    m0, p0 = orc.stft_transform(hs, 20, 5)
    assert (m1 - m0).abs().max().item() < 5e-6
    assert (wrap_phase_err(p1, p0) * m0.double()).max().item() < 5e-6


def test_explicit_linear_resample_is_bit_exact_vs_aten():
    """the fp32 rounding points the CUDA source kernel mirrors (fma(ratio, i+.5, -.5), fma(l0,x0,l1*x1))"""
    import torch.nn.functional as F
    T = 200
    inp = synth.synth_inputs(CFGS["istftnet"], 1, T, with_noise=False)
    harm = torch.arange(1, 10, dtype=torch.float32)
    rad = ((inp["F0_curve"][:, :, None] * harm) / 24000) % 1
    pt = (torch.cumsum(rad, dim=1) * 2 * np.pi).transpose(1, 2) * 300
    up = F.interpolate(pt, scale_factor=300, mode="linear").numpy()
    ex = orc.linear_resample_explicit(pt.contiguous().numpy(), 300 * T, np.float32(1.0 / 300.0))
    assert np.array_equal(up, ex)
    cs = np.cumsum(rad.numpy().astype(np.float64), axis=1).astype(np.float32)  # ATen CPU: double accumulate
    assert np.array_equal((cs * np.float32(2) * np.float32(np.pi)).astype(np.float32),
                          (torch.cumsum(rad, dim=1) * 2 * np.pi).numpy())


def test_stft_istft_roundtrip_and_envelope():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 600, generator=g, dtype=torch.float64)
    mag, ph = orc.stft_transform(x, 20, 5)
    assert mag.shape == (2, 11, 121)
    y = orc.istft_inverse(mag, ph, 20, 5)
    assert y.shape == (2, 1, 600)
    assert (y[:, 0] - x).abs().max().item() < 1e-9  # hann/hop-5 satisfies NOLA: exact reconstruction


def test_fp64_oracle_close_to_fp32_oracle_with_har():
    g = gold("istftnet", 8)
    a = orc.decoder_forward(sd_of("istftnet"), CFGS["istftnet"], g["asr"], g["F0_curve"], g["N"], g["s"],
                            har=g["tap_har"], dtype=torch.float64)
    assert (a.float() - g["wav"]).abs().max().item() < 1e-4


def test_metrics():
    g = torch.Generator().manual_seed(0)
    a = torch.randn(1, 1, 24000, generator=g)
    assert orc.snr_db(a, a + 0.01 * torch.randn(1, 1, 24000, generator=g)) == pytest.approx(40, abs=0.5)
    assert orc.log_mel_l1(a, a) == 0.0
    assert orc.log_mel(a).shape == (1, 80, 81)


# ------------------------------------------------------------- F0/N predictor
def test_f0n_spec_matches_reference_keys():
    ref = json.load(open(os.path.join(GOLD, "keys_f0n.json")))
    mine = {n: list(s) for n, s, *_ in synth.f0n_param_spec(synth.F0N_CFG)}
    assert mine == ref and len(ref) == 82


@pytest.mark.parametrize("T", [20, 51])
def test_f0n_oracle_vs_reference(T):
    """oracle/predictor_oracle.py against ProsodyPredictor.F0Ntrain of the real reference
    (oracle/make_golden_f0n.py): LSTM output and both curves"""
    from oracle import predictor_oracle as porc
    g = np.load(os.path.join(GOLD, f"f0n_T{T}.npz"))
    g = {k: torch.from_numpy(g[k]) for k in g.files}
    sd = synth.synth_f0n_state_dict(synth.F0N_CFG, seed=0)
    lstm = porc.shared_lstm(g["x"], porc.Params(sd))
    assert (lstm - g["lstm"]).abs().max().item() < 2e-6
    F0a, Na = porc.f0n_forward(sd, g["x"], g["s"], aten_lstm=True)   # the timed baseline's nn.LSTM route
    assert (F0a - g["F0"]).abs().max().item() < 5e-6 and (Na - g["N"]).abs().max().item() < 5e-6
    F0, N = porc.f0n_forward(sd, g["x"], g["s"])
    assert F0.shape == g["F0"].shape == (g["x"].shape[0], 2 * T)
    assert (F0 - g["F0"]).abs().max().item() < 5e-6
    assert (N - g["N"]).abs().max().item() < 5e-6


# ------------------------------------------------------------- style sampler
@pytest.mark.parametrize("variant", ["style", "lj"])
def test_sampler_spec_matches_reference_keys(variant):
    cfg = synth.SAMPLER_CFG if variant == "style" else synth.SAMPLER_CFG_LJ
    ref = json.load(open(os.path.join(GOLD, f"keys_sampler_{variant}.json")))
    mine = {n: list(s) for n, s, *_ in synth.sampler_param_spec(cfg)}
    assert mine == ref and len(ref) == (48 if variant == "style" else 46)


@pytest.mark.parametrize("name", ["style", "style_cfg", "lj"])
def test_sampler_oracle_vs_reference(name):
    """oracle/sampler_oracle.py against DiffusionSampler(ADPM2Sampler, KarrasSchedule) and KDiffusion.denoise_fn of
    the real reference (oracle/make_golden_sampler.py), with and without classifier-free guidance, both networks"""
    from oracle import sampler_oracle as sorc
    cfg = synth.SAMPLER_CFG_LJ if name == "lj" else synth.SAMPLER_CFG
    g = np.load(os.path.join(GOLD, f"sampler_{name}.npz"))
    g = {k: torch.from_numpy(np.asarray(g[k])) for k in g.files}
    sd = synth.synth_sampler_state_dict(cfg, seed=0)
    feat, scale, steps = g.get("features"), float(g["embedding_scale"]), int(g["num_steps"])
    out = sorc.sample(sd, cfg, g["noise"], g["embedding"], feat, steps, scale, g["step_noise"])
    assert out.shape == g["out"].shape and (out - g["out"]).abs().max().item() < 5e-6
    den = sorc.denoise(sorc.Net(sd, cfg), g["noise"] * 0.7, 0.45, 0.2, embedding=g["embedding"], features=feat,
                       embedding_scale=scale)
    assert (den - g["denoised"]).abs().max().item() < 5e-6


def test_sampler_schedule_known_values():
    """KarrasSchedule(1e-4, 3.0, rho=9) at 5 steps: endpoints exact, monotone, trailing 0; ADPM2 sigmas satisfy
    sigma_up^2 + sigma_down^2 = sigma_next^2 and sigma_down <= sigma_mid <= sigma"""
    from oracle import sampler_oracle as sorc
    s = sorc.karras_sigmas(5, 1e-4, 3.0, 9.0)
    assert s.shape == (6,) and abs(s[0].item() - 3.0) < 1e-5 and abs(s[4].item() - 1e-4) < 1e-8 and s[5].item() == 0.0
    assert (s[:-1][1:] < s[:-1][:-1]).all()
    up, down, mid = sorc.adpm2_sigmas(s[1], s[2])
    assert abs(up ** 2 + down ** 2 - float(s[2]) ** 2) < 1e-6 and down <= float(mid) <= float(s[1])


@pytest.mark.parametrize("typ", ["istftnet", "hifigan"])
def test_alignment_oracle_vs_notebook_cell(typ):
    """oracle/alignment_oracle.py against the statements of Demo/Inference_LibriTTS.ipynb cell 16 themselves
    (oracle/make_golden_alignment.py cuts them out of the notebook JSON and exec()s them): bit exact, incl. a
    one-token text and the ``pred_dur[-1] = 1`` branch.  This is what pins row N3."""
    from oracle import alignment_oracle as aorc
    g = np.load(os.path.join(GOLD, "alignment_notebook.npz"))
    for name in "abcd":
        dur = torch.from_numpy(g[f"{name}_dur"])
        frames = int(dur.sum())
        aln = aorc.one_hot_alignment(dur, frames)
        assert torch.equal(aln, torch.from_numpy(g[f"{name}_{typ}_aln"]))
        d, t_en = torch.from_numpy(g[f"{name}_d"]), torch.from_numpy(g[f"{name}_t_en"])
        en = aorc.expand(d.transpose(-1, -2), dur[None], frames, typ == "hifigan")
        asr = aorc.expand(t_en, dur[None], frames, typ == "hifigan")
        assert torch.equal(en, torch.from_numpy(g[f"{name}_{typ}_en"]))
        assert torch.equal(asr, torch.from_numpy(g[f"{name}_{typ}_asr"]))


def test_log_mel_oracle_vs_reference_front_end():
    """oracle log_mel against meldataset.py:35-43 executed verbatim with torchaudio (oracle/make_golden_logmel.py);
    the reference computes in fp32, the oracle in double"""
    g = np.load(os.path.join(GOLD, "logmel.npz"))
    for n in "abc":
        m = orc.log_mel(torch.from_numpy(g[f"{n}_wave"]))
        ref = torch.from_numpy(g[f"{n}_mel"]).double()
        assert m.shape == ref.shape and (m - ref).abs().max().item() < 1e-4

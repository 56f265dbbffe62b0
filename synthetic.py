"""Synthetic weights and inputs for the StyleTTS2 waveform-decoder path.

Neutral data generator: it contains no decoder arithmetic.  ``bench.py``, ``tests/``,
``__graft_entry__.smoke()`` and ``oracle/make_golden.py`` use it to build the same random-init
state dict and seeded inputs on both sides of a comparison; the product package
(``styletts2_b200``) does not import it.

Why synthetic weights: the reference ships no checkpoints
(``/root/reference/.MISSING_LARGE_BLOBS``) and BASELINE.json asks for random-init
weights.  The reference's own random init depends on the torch RNG *call order* of
its constructors, which we do not want to mirror line by line; instead every tensor
is drawn from a generator seeded by ``crc32(name) ^ seed`` so the same state dict
can be rebuilt anywhere (here, on the GPU box) from the key/shape list alone.

Two regimes (``init=``):

``"reference"`` (default; what bench.py, smoke() and the main goldens use) — the distributions the
reference's constructors REALLY leave behind under ``torch.manual_seed(0); Decoder(...)``:

* every conv / linear weight (``weight_v`` of the weight-normed ones included) and bias:
  torch's default kaiming-uniform U(-1/sqrt(fan_in), 1/sqrt(fan_in)).  ``init_weights``
  (``Modules/utils.py:1-4`` applied at ``istftnet.py:38,48,343-344``) does NOT change this: it
  writes ``m.weight.data.normal_()`` on an old-style weight-normed module, where ``.weight`` is an
  attribute recomputed from ``weight_g`` / ``weight_v`` by a forward pre-hook, so the N(0, 0.01)
  draw is overwritten on the first forward and ``weight_v`` keeps its kaiming-uniform values
  (measured: ``resblocks.3.convs1.0.weight_v`` std 0.0295 = 0.051/sqrt(3), not 0.01).
* ``weight_g`` = ``||v||`` exactly (what ``torch.nn.utils.weight_norm`` sets at wrap time).
* Snake ``alpha`` = 1 (``istftnet.py:62-63``).

The per-layer statistics of this regime are pinned against the real reference by
``tests/golden/init_stats_*.json`` (``oracle/make_golden.py``, ``tests/test_oracle_golden.py``).

``"legacy"`` (round-1 goldens, kept as a second regime because it exercises things the reference
init does not): ``init_weights`` layers N(0, 0.01), ``weight_g`` = ``||v||`` x (1 +- 10 %) so that
weight-norm folding is not the identity, Snake ``alpha`` = 1 +- 20 % per channel.

The key set / shapes produced by :func:`param_spec` are pinned against the real
reference modules by ``oracle/make_golden.py`` (-> ``tests/golden/keys_*.json``).
"""
from __future__ import annotations

import math
import zlib
from typing import Dict, List, Tuple

import numpy as np
import torch

ISTFTNET_CFG = dict(
    kind="istftnet", dim_in=512, style_dim=128, dim_out=80,
    resblock_kernel_sizes=[3, 7, 11], upsample_rates=[10, 6],
    upsample_initial_channel=512,
    resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    upsample_kernel_sizes=[20, 12], gen_istft_n_fft=20, gen_istft_hop_size=5,
)  # /root/reference/Configs/config.yml:51-60 + models.py:1008-1019

HIFIGAN_CFG = dict(
    kind="hifigan", dim_in=512, style_dim=128, dim_out=80,
    resblock_kernel_sizes=[3, 7, 11], upsample_rates=[10, 5, 3, 2],
    upsample_initial_channel=512,
    resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    upsample_kernel_sizes=[20, 10, 6, 4],
)  # /root/reference/Configs/config_libritts.yml:49-56 + models.py:1021-1032


def samples_per_frame(cfg) -> int:
    u = int(np.prod(cfg["upsample_rates"]))
    if cfg["kind"] == "istftnet":
        u *= cfg["gen_istft_hop_size"]
    return u


# --------------------------------------------------------------------------- spec
# entry: (name, shape, kind, fan_in)   kind in {v_default, v_init, g, bias, w, alpha}
Spec = List[Tuple[str, Tuple[int, ...], str, int]]


def _wn_conv(spec: Spec, p: str, cout: int, cin_g: int, k: int, init: bool, bias=True,
             transposed_cin: int | None = None):
    """weight-normed Conv1d (or ConvTranspose1d when transposed_cin is given: then
    weight_v is [Cin, Cout/groups, k] and g is per *input* channel, dim 0)."""
    fan_in = cin_g * k
    if transposed_cin is None:
        vshape, gshape = (cout, cin_g, k), (cout, 1, 1)
    else:
        vshape, gshape = (transposed_cin, cin_g, k), (transposed_cin, 1, 1)
        fan_in = cin_g * k  # torch: fan_in of ConvTranspose weight = size(1)*k
    if bias:
        spec.append((p + ".bias", (cout,), "bias", fan_in))
    spec.append((p + ".weight_g", gshape, "g", fan_in))
    spec.append((p + ".weight_v", vshape, "v_init" if init else "v_default", fan_in))


def _adain(spec: Spec, p: str, style: int, c: int):
    spec.append((p + ".fc.weight", (2 * c, style), "w", style))
    spec.append((p + ".fc.bias", (2 * c,), "bias", style))


def _adain_resblk1d(spec: Spec, p: str, din: int, dout: int, style: int, upsample: bool):
    # istftnet.py:409-432
    _wn_conv(spec, p + ".conv1", dout, din, 3, False)
    _wn_conv(spec, p + ".conv2", dout, dout, 3, False)
    _adain(spec, p + ".norm1", style, din)
    _adain(spec, p + ".norm2", style, dout)
    if din != dout:
        _wn_conv(spec, p + ".conv1x1", dout, din, 1, False, bias=False)
    if upsample:
        # depthwise ConvTranspose1d(din, din, 3, stride 2, groups=din): weight [din,1,3]
        _wn_conv(spec, p + ".pool", din, 1, 3, False, transposed_cin=din)


def _adain_resblock1(spec: Spec, p: str, c: int, k: int, style: int):
    # istftnet.py:27-63
    for grp in ("convs1", "convs2"):
        for i in range(3):
            _wn_conv(spec, f"{p}.{grp}.{i}", c, c, k, True)
    for grp in ("adain1", "adain2"):
        for i in range(3):
            _adain(spec, f"{p}.{grp}.{i}", style, c)
    for grp in ("alpha1", "alpha2"):
        for i in range(3):
            spec.append((f"{p}.{grp}.{i}", (1, c, 1), "alpha", 1))


def param_spec(cfg) -> Spec:
    """Names/shapes of ``Decoder.state_dict()`` (istftnet.py:466-496, hifigan.py:415-442)."""
    spec: Spec = []
    style = cfg["style_dim"]
    din = cfg["dim_in"]
    for i in range(3):
        _adain_resblk1d(spec, f"decode.{i}", 1024 + 2 + 64, 1024, style, False)
    _adain_resblk1d(spec, "decode.3", 1024 + 2 + 64, 512, style, True)
    _adain_resblk1d(spec, "encode", din + 2, 1024, style, False)
    _wn_conv(spec, "F0_conv", 1, 1, 3, False)
    _wn_conv(spec, "N_conv", 1, 1, 3, False)
    _wn_conv(spec, "asr_res.0", 64, 512, 1, False)

    g = "generator"
    spec.append((g + ".m_source.l_linear.weight", (1, 9), "w", 9))
    spec.append((g + ".m_source.l_linear.bias", (1,), "bias", 9))
    rates, ks = cfg["upsample_rates"], cfg["upsample_kernel_sizes"]
    c0 = cfg["upsample_initial_channel"]
    n_up = len(rates)
    istft = cfg["kind"] == "istftnet"
    src_ch = cfg["gen_istft_n_fft"] + 2 if istft else 1
    for i, (u, k) in enumerate(zip(rates, ks)):
        cin, cout = c0 // (2 ** i), c0 // (2 ** (i + 1))
        # ConvTranspose1d(cin, cout, k, u): weight_v [cin, cout, k]; torch fan_in = cout*k
        _wn_conv(spec, f"{g}.ups.{i}", cout, cout, k, True, transposed_cin=cin)
        if i + 1 < n_up:
            stride_f0 = int(np.prod(rates[i + 1:]))
            kk = stride_f0 * 2
            rk = 7
        else:
            kk, rk = 1, 11
        spec.append((f"{g}.noise_convs.{i}.weight", (cout, src_ch, kk), "w", src_ch * kk))
        spec.append((f"{g}.noise_convs.{i}.bias", (cout,), "bias", src_ch * kk))
        _adain_resblock1(spec, f"{g}.noise_res.{i}", cout, rk, style)
        for j, k_r in enumerate(cfg["resblock_kernel_sizes"]):
            _adain_resblock1(spec, f"{g}.resblocks.{i * len(cfg['resblock_kernel_sizes']) + j}",
                             cout, k_r, style)
    ch_last = c0 // (2 ** n_up)
    if istft:
        _wn_conv(spec, g + ".conv_post", cfg["gen_istft_n_fft"] + 2, ch_last, 7, True)
    else:
        _wn_conv(spec, g + ".conv_post", 1, ch_last, 7, True)
        spec.append((g + ".alphas.0", (1, c0, 1), "alpha", 1))
        for i in range(n_up):
            spec.append((f"{g}.alphas.{i + 1}", (1, c0 // (2 ** (i + 1)), 1), "alpha", 1))
    return spec


# ------------------------------------------------------------------ synthetic draw
def _gen_for(name: str, seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(name.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    return g


def synth_state_dict(cfg, seed: int = 0, init: str = "reference") -> Dict[str, torch.Tensor]:
    """Deterministic fp32 state dict with the reference's key set (see module doc)."""
    assert init in ("reference", "legacy")
    legacy = init == "legacy"
    spec = param_spec(cfg)
    sd: Dict[str, torch.Tensor] = {}
    for name, shape, kind, fan_in in spec:
        g = _gen_for(name, seed)
        if kind == "g":
            continue  # derived from v below
        if kind == "v_init" and legacy:
            t = torch.randn(shape, generator=g) * 0.01
        elif kind in ("v_init", "v_default", "w", "bias"):
            bound = 1.0 / math.sqrt(max(fan_in, 1))
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif kind == "alpha":
            t = 1.0 + 0.2 * (torch.rand(shape, generator=g) * 2 - 1) if legacy else torch.ones(shape)
        else:
            raise ValueError(kind)
        sd[name] = t.float().contiguous()
    for name, shape, kind, _ in spec:
        if kind != "g":
            continue
        v = sd[name[:-1] + "v"]
        nrm = v.reshape(v.shape[0], -1).norm(dim=1).reshape(shape)
        pert = 1.0 + 0.1 * (torch.rand(shape, generator=_gen_for(name, seed)) * 2 - 1) if legacy else 1.0
        sd[name] = (nrm * pert).float().contiguous()
    return {name: sd[name] for name, *_ in spec}


def init_stats(sd: Dict[str, torch.Tensor]) -> Dict[str, List[float]]:
    """per-tensor [std, abs-max, mean] of a decoder state dict; for ``weight_g`` the std / max / mean of
    ``g / ||v||`` instead (1 at the reference's init).  Used to pin :func:`synth_state_dict` against the
    reference's own ``torch.manual_seed(0); Decoder(...)`` (tests/golden/init_stats_*.json)."""
    out = {}
    for k, t in sd.items():
        t = t.detach().double()
        if k.endswith("weight_g"):
            v = sd[k[:-1] + "v"].detach().double()
            t = t.flatten() / v.reshape(v.shape[0], -1).norm(dim=1)
        std = float(t.std()) if t.numel() > 1 else 0.0
        out[k] = [std, float(t.abs().max()), float(t.mean())]
    return out


def synth_inputs(cfg, B: int, T: int, seed: int = 1234, with_noise: bool = True):
    """Seeded decoder inputs: asr ~ N(0,1) [B,512,T/2]; F0 a smooth
    70-400 Hz contour with ~30 % unvoiced (zero) frames in runs of 5-40; N ~ N(0,1);
    s ~ N(0,1) one row per item; noise = the ``randn_like(sine_waves)`` draw of
    SineGen.forward (istftnet.py:241), shape [B, T*spf, 9]."""
    assert T % 2 == 0
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    asr = torch.randn(B, cfg["dim_in"], T // 2, generator=g)
    t = torch.arange(T, dtype=torch.float32)[None, :]
    phi = torch.rand(B, 1, generator=g) * 2 * math.pi
    f0 = 180 + 60 * torch.sin(2 * math.pi * t / 97 + phi) + 20 * torch.randn(B, T, generator=g)
    f0 = f0.clamp(70, 400)
    rs = np.random.RandomState(seed)
    for b in range(B):
        n_zero, guard = 0, 0
        while n_zero < 0.3 * T and guard < 1000:
            guard += 1
            run = int(rs.randint(5, 41)) if T >= 160 else int(rs.randint(1, max(T // 8, 2) + 1))
            st = int(rs.randint(0, max(T - run, 1)))
            f0[b, st:st + run] = 0.0
            n_zero = int((f0[b] == 0).sum())
    N = torch.randn(B, T, generator=g)
    s = torch.randn(B, cfg["style_dim"], generator=g)
    out = dict(asr=asr.contiguous(), F0_curve=f0.contiguous(), N=N.contiguous(), s=s.contiguous())
    if with_noise:
        out["noise"] = torch.randn(B, T * samples_per_frame(cfg), 9, generator=g)
    return out


# ------------------------------------------------------------- F0/N predictor
F0N_CFG = {"style_dim": 128, "d_hid": 512}


def f0n_param_spec(cfg) -> Spec:
    """Names/shapes of the F0Ntrain part of ``ProsodyPredictor.state_dict()`` (models.py:562-590):
    ``shared`` (bidirectional LSTM), ``F0.{0,1,2}``, ``N.{0,1,2}``, ``F0_proj``, ``N_proj``."""
    spec: Spec = []
    style, d = cfg["style_dim"], cfg["d_hid"]
    hid = d // 2
    for sfx in ("", "_reverse"):
        spec.append((f"shared.weight_ih_l0{sfx}", (4 * hid, d + style), "w", hid))
        spec.append((f"shared.weight_hh_l0{sfx}", (4 * hid, hid), "w", hid))
        spec.append((f"shared.bias_ih_l0{sfx}", (4 * hid,), "bias", hid))
        spec.append((f"shared.bias_hh_l0{sfx}", (4 * hid,), "bias", hid))
    for br in ("F0", "N"):
        _adain_resblk1d(spec, f"{br}.0", d, d, style, False)
        _adain_resblk1d(spec, f"{br}.1", d, hid, style, True)
        _adain_resblk1d(spec, f"{br}.2", hid, hid, style, False)
        spec.append((f"{br}_proj.weight", (1, hid, 1), "w", hid))
        spec.append((f"{br}_proj.bias", (1,), "bias", hid))
    return spec


def synth_f0n_state_dict(cfg, seed: int = 0) -> Dict[str, torch.Tensor]:
    spec = f0n_param_spec(cfg)
    sd: Dict[str, torch.Tensor] = {}
    for name, shape, kind, fan_in in spec:
        if kind == "g":
            continue
        g = _gen_for("f0n." + name, seed)
        bound = 1.0 / math.sqrt(max(fan_in, 1))
        sd[name] = ((torch.rand(shape, generator=g) * 2 - 1) * bound).float().contiguous()
    for name, shape, kind, _ in spec:
        if kind != "g":
            continue
        v = sd[name[:-1] + "v"]
        nrm = v.reshape(v.shape[0], -1).norm(dim=1).reshape(shape)
        pert = 1.0 + 0.1 * (torch.rand(shape, generator=_gen_for("f0n." + name, seed)) * 2 - 1)
        sd[name] = (nrm * pert).float().contiguous()
    return {name: sd[name] for name, *_ in spec}


def synth_f0n_inputs(cfg, B: int, T: int, seed: int = 77):
    """``en`` [B, d_hid + style, T] (duration-expanded prosody encoding) and ``s`` [B, style]."""
    g = torch.Generator().manual_seed(seed)
    return {"x": torch.randn(B, cfg["d_hid"] + cfg["style_dim"], T, generator=g),
            "s": torch.randn(B, cfg["style_dim"], generator=g)}


# ---------------------------------------------------------------------------------------------
# Style sampler: StyleTransformer1d / Transformer1d denoiser of the ADPM2 sampler
# ---------------------------------------------------------------------------------------------
# models.py:1057-1069 with Configs/config_libritts.yml:68-75 (style_dim 128 -> channels 256, PL-BERT
# hidden size 768): multispeaker = StyleTransformer1d(context_features=256), else Transformer1d.
SAMPLER_CFG = {"channels": 256, "context_features": 256, "context_embedding_features": 768, "num_layers": 3,
               "num_heads": 8, "head_features": 64, "multiplier": 2, "embedding_max_length": 512}
SAMPLER_CFG_LJ = dict(SAMPLER_CFG, context_features=0)  # Transformer1d (Configs/config.yml, multispeaker: false)


def sampler_param_spec(cfg) -> Spec:
    """Names/shapes of ``StyleTransformer1d.state_dict()`` (context_features > 0) or
    ``Transformer1d.state_dict()`` (Modules/diffusion/modules.py:40-119, 280-355)."""
    spec: Spec = []
    C, S, E = cfg["channels"], cfg["context_features"], cfg["context_embedding_features"]
    F_, mid = C + E, cfg["num_heads"] * cfg["head_features"]
    for i in range(cfg["num_layers"]):
        p = f"blocks.{i}"
        for nm in ("norm", "norm_context"):
            if S:
                spec.append((f"{p}.attention.{nm}.fc.weight", (2 * F_, S), "w", S))
                spec.append((f"{p}.attention.{nm}.fc.bias", (2 * F_,), "bias", S))
            else:
                spec.append((f"{p}.attention.{nm}.weight", (F_,), "ln_w", 1))
                spec.append((f"{p}.attention.{nm}.bias", (F_,), "ln_b", 1))
        spec.append((f"{p}.attention.to_q.weight", (mid, F_), "w", F_))
        spec.append((f"{p}.attention.to_kv.weight", (2 * mid, F_), "w", F_))
        spec.append((f"{p}.attention.attention.to_out.weight", (F_, mid), "w", mid))
        spec.append((f"{p}.attention.attention.to_out.bias", (F_,), "bias", mid))
        spec.append((f"{p}.feed_forward.0.weight", (F_ * cfg["multiplier"], F_), "w", F_))
        spec.append((f"{p}.feed_forward.0.bias", (F_ * cfg["multiplier"],), "bias", F_))
        spec.append((f"{p}.feed_forward.2.weight", (F_, F_ * cfg["multiplier"]), "w", F_ * cfg["multiplier"]))
        spec.append((f"{p}.feed_forward.2.bias", (F_,), "bias", F_ * cfg["multiplier"]))
    spec.append(("to_out.1.weight", (C, F_, 1), "w", F_))
    spec.append(("to_out.1.bias", (C,), "bias", F_))
    for j in (0, 2):
        spec.append((f"to_mapping.{j}.weight", (F_, F_), "w", F_))
        spec.append((f"to_mapping.{j}.bias", (F_,), "bias", F_))
    spec.append(("to_time.0.0.weights", (C // 2,), "randn", 1))
    spec.append(("to_time.0.1.weight", (F_, C + 1), "w", C + 1))
    spec.append(("to_time.0.1.bias", (F_,), "bias", C + 1))
    if S:
        spec.append(("to_features.0.weight", (F_, S), "w", S))
        spec.append(("to_features.0.bias", (F_,), "bias", S))
    spec.append(("fixed_embedding.embedding.weight", (cfg["embedding_max_length"], E), "randn", 1))
    return spec


def synth_sampler_state_dict(cfg, seed: int = 0) -> Dict[str, torch.Tensor]:
    """torch-default-like init (U(+-1/sqrt(fan_in)) for linears, N(0,1) for the Fourier weights and the
    embedding table, LayerNorm weights around 1), each tensor from its own seeded generator"""
    sd: Dict[str, torch.Tensor] = {}
    for name, shape, kind, fan_in in sampler_param_spec(cfg):
        g = _gen_for("sampler." + name, seed)
        if kind == "randn":
            sd[name] = torch.randn(shape, generator=g).float()
        elif kind == "ln_w":
            sd[name] = (1.0 + 0.2 * torch.randn(shape, generator=g)).float()
        elif kind == "ln_b":
            sd[name] = (0.2 * torch.randn(shape, generator=g)).float()
        else:
            bound = 1.0 / math.sqrt(max(fan_in, 1))
            sd[name] = ((torch.rand(shape, generator=g) * 2 - 1) * bound).float().contiguous()
    return sd


def synth_sampler_inputs(cfg, B: int, n_tokens: int, num_steps: int = 5, seed: int = 5):
    """noise [B, 1, channels], embedding (PL-BERT output stand-in) [B, n, E], features (ref_s) [B, S],
    step_noise [num_steps - 1, B, 1, channels]: the randn_like draws of ADPM2Sampler.step (sampler.py:510)"""
    g = torch.Generator().manual_seed(seed)
    C, S, E = cfg["channels"], cfg["context_features"], cfg["context_embedding_features"]
    out = {"noise": torch.randn(B, 1, C, generator=g), "embedding": torch.randn(B, n_tokens, E, generator=g),
           "step_noise": torch.randn(max(num_steps - 1, 1), B, 1, C, generator=g)}
    out["features"] = torch.randn(B, S, generator=g) if S else None
    return out

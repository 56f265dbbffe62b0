"""CPU oracle: a functional restatement of StyleTTS2's waveform decoder forward pass.

TEST INFRASTRUCTURE ONLY.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import this file;
the product (``styletts2_b200``) never does and fails loudly without its CUDA library.

What is restated (all ``file:line`` are under /root/reference):
  * ``Decoder.forward``            Modules/istftnet.py:498-527 (hifigan.py:445-474)
  * ``AdainResBlk1d``              Modules/istftnet.py:409-453
  * ``AdaIN1d``                    Modules/istftnet.py:15-25
  * ``AdaINResBlock1`` (+Snake)    Modules/istftnet.py:27-75
  * ``Generator.forward``          Modules/istftnet.py:349-379, Modules/hifigan.py:318-346
  * ``SineGen`` / ``SourceModuleHnNSF``  Modules/istftnet.py:146-186,218-246,282-296
  * ``TorchSTFT.transform/inverse``      Modules/istftnet.py:91-104
  * old-style ``weight_norm``            Modules/istftnet.py:5 (w = g * v / ||v||, dim 0)

Third-party arithmetic: the dense contractions (``conv1d``, ``conv_transpose1d``,
``linear``) live in PyTorch/ATen (the reference pins no version, requirements.txt:4;
this image has torch 2.11.0), which is present both here and on the GPU box, so the
oracle calls ``torch.nn.functional`` for them on the CPU.  Everything whose behaviour is
*not* obvious from the call (instance-norm statistics, Snake, the harmonic source and
its two linear resamplings, the 20-point STFT / inverse STFT with overlap-add and
envelope division, reflection pad, polyphase structure) is written out explicitly.

Pinning: the reference holds no tests or golden vectors, so parity
is pinned by ``oracle/make_golden.py``, which imports the reference's own modules in
the build container, loads the synthetic state dict from ``synthetic.py`` and stores
inputs/outputs under ``tests/golden/``.  ``tests/test_oracle_golden.py`` checks this
file against those vectors.

Every function takes a ``dtype`` (float32 mirrors the reference, float64 gives a
numerically cleaner yardstick) and works on NCL tensors like the reference.

Library mode (``decoder_forward(..., library=True)``): the three places where this file writes the
arithmetic out (instance norm, STFT, inverse STFT) call the SAME torch entry points the reference calls
instead (``F.instance_norm`` = nn.InstanceNorm1d, ``torch.stft`` / ``torch.istft`` as in
istftnet.py:91-104), and every tensor follows the device of the inputs.  That is the reference's own op
sequence on ATen; ``bench.py --impl reference-gpu`` / the ``library_baseline`` block run it on the B200
(cuDNN / cuFFT, eager fp32 with TF32 off as train_second.py:183-186, and under bf16 autocast) as the
library baseline the kernels have to beat.  ``tests/test_oracle_golden.py`` checks library mode against
the same goldens of the real reference.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

LRELU_SLOPE = 0.1  # istftnet.py:13
SAMPLING_RATE = 24000  # istftnet.py:310
VOICED_THRESHOLD = 10.0  # istftnet.py:312
SINE_AMP = 0.1  # istftnet.py:128
NOISE_STD = 0.003  # istftnet.py:128
N_HARM = 9  # harmonic_num=8 -> dim 9, istftnet.py:135,312


# ------------------------------------------------------------------- primitives
def fold_weight_norm(g: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """torch.nn.utils.weight_norm, dim=0: w = g * v / ||v||_2 over all dims but 0."""
    nrm = v.reshape(v.shape[0], -1).norm(dim=1).reshape([-1] + [1] * (v.dim() - 1))
    return v * (g / nrm)


class Params:
    """state dict view with weight-norm folded on access, in a working dtype."""

    def __init__(self, sd: Dict[str, torch.Tensor], dtype=torch.float32, library: bool = False):
        self.sd = {(k[7:] if k.startswith("module.") else k): v for k, v in sd.items()}
        self.dtype = dtype
        self.library = library   # see the module doc: torch's own instance_norm / stft / istft
        self._cache: Dict[str, torch.Tensor] = {}

    def get(self, name: str) -> torch.Tensor:
        return self.sd[name].to(self.dtype)

    def has(self, name: str) -> bool:
        return name in self.sd

    def weight(self, prefix: str) -> torch.Tensor:
        if prefix not in self._cache:
            if prefix + ".weight_v" in self.sd:
                # fold in fp64 when the working dtype is fp64, else like torch (fp32)
                w = fold_weight_norm(self.get(prefix + ".weight_g"), self.get(prefix + ".weight_v"))
            else:
                w = self.get(prefix + ".weight")
            self._cache[prefix] = w
        return self._cache[prefix]

    def bias(self, prefix: str) -> Optional[torch.Tensor]:
        return self.get(prefix + ".bias") if prefix + ".bias" in self.sd else None


def instance_norm(x: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    """nn.InstanceNorm1d(affine=False): per-(b,c) mean / biased variance over time."""
    mean = x.mean(dim=2, keepdim=True)
    var = ((x - mean) ** 2).mean(dim=2, keepdim=True)
    return (x - mean) / torch.sqrt(var + eps)


def adain(x, s, P: Params, prefix: str):
    """AdaIN1d.forward, istftnet.py:21-25."""
    h = F.linear(s, P.get(prefix + ".fc.weight"), P.get(prefix + ".fc.bias"))
    gamma, beta = torch.chunk(h[:, :, None], 2, dim=1)
    xn = F.instance_norm(x, eps=1e-5) if P.library else instance_norm(x)
    return (1 + gamma) * xn + beta


def snake(x, alpha):
    """Snake1D as written at istftnet.py:69: x + (1/a) * sin(a x)^2."""
    return x + (1 / alpha) * (torch.sin(alpha * x) ** 2)


def get_padding(k: int, d: int = 1) -> int:
    return int((k * d - d) / 2)  # Modules/utils.py:6-7


def adain_resblock1(x, s, P: Params, prefix: str, k: int, dilations=(1, 3, 5)):
    """AdaINResBlock1.forward, istftnet.py:66-75."""
    for i, d in enumerate(dilations):
        xt = adain(x, s, P, f"{prefix}.adain1.{i}")
        xt = snake(xt, P.get(f"{prefix}.alpha1.{i}"))
        xt = F.conv1d(xt, P.weight(f"{prefix}.convs1.{i}"), P.bias(f"{prefix}.convs1.{i}"),
                      dilation=d, padding=get_padding(k, d))
        xt = adain(xt, s, P, f"{prefix}.adain2.{i}")
        xt = snake(xt, P.get(f"{prefix}.alpha2.{i}"))
        xt = F.conv1d(xt, P.weight(f"{prefix}.convs2.{i}"), P.bias(f"{prefix}.convs2.{i}"),
                      dilation=1, padding=get_padding(k, 1))
        x = xt + x
    return x


def adain_resblk1d(x, s, P: Params, prefix: str, upsample: bool):
    """AdainResBlk1d.forward, istftnet.py:434-453 (dropout p=0 is a no-op)."""
    learned_sc = P.has(prefix + ".conv1x1.weight_v")
    # residual
    r = adain(x, s, P, prefix + ".norm1")
    r = F.leaky_relu(r, 0.2)
    if upsample:  # weight-normed depthwise ConvTranspose1d(k3,s2,p1,op1), istftnet.py:423
        r = F.conv_transpose1d(r, P.weight(prefix + ".pool"), P.bias(prefix + ".pool"),
                               stride=2, padding=1, output_padding=1, groups=r.shape[1])
    r = F.conv1d(r, P.weight(prefix + ".conv1"), P.bias(prefix + ".conv1"), padding=1)
    r = adain(r, s, P, prefix + ".norm2")
    r = F.leaky_relu(r, 0.2)
    r = F.conv1d(r, P.weight(prefix + ".conv2"), P.bias(prefix + ".conv2"), padding=1)
    # shortcut
    sc = x
    if upsample:  # UpSample1d: nearest x2, istftnet.py:460-464
        sc = sc.repeat_interleave(2, dim=2)
    if learned_sc:
        sc = F.conv1d(sc, P.weight(prefix + ".conv1x1"))
    return (r + sc) / math.sqrt(2)


# -------------------------------------------------------------- harmonic source
def _fma32(a, b, c):
    """float32 fused multiply-add emulated in float64 (product of two f32 is exact in f64)."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64)
            + np.asarray(c, np.float64)).astype(np.float32)


def linear_resample_explicit(x: np.ndarray, out_len: int, ratio_f32: np.float32) -> np.ndarray:
    """ATen CPU ``upsample_linear1d`` (align_corners=False) written out in float32 with the
    exact rounding points of this torch build [probed: bit-identical on 1.08 M samples]:
    src = fma(ratio, i+0.5, -0.5) clamped at 0; i0 = floor(src); i1 = min(i0+1, L-1);
    l1 = src - i0; l0 = 1 - l1; out = fma(l0, x[i0], l1*x[i1]).   x: [..., L] float32."""
    L = x.shape[-1]
    i = np.arange(out_len, dtype=np.float32)
    src = _fma32(np.full_like(i, ratio_f32), i + np.float32(0.5), np.full_like(i, -0.5))
    src = np.maximum(src, np.float32(0.0)).astype(np.float32)
    i0 = np.minimum(src.astype(np.int64), L - 1)
    i1 = np.minimum(i0 + 1, L - 1)
    l1 = (src - i0.astype(np.float32)).astype(np.float32)
    l0 = (np.float32(1.0) - l1).astype(np.float32)
    x0, x1 = x[..., i0], x[..., i1]
    return _fma32(l0 * np.ones_like(x0), x0, (l1 * x1).astype(np.float32))


def sine_gen(f0_frames: torch.Tensor, noise: torch.Tensor, spf: int, dtype=torch.float32):
    """SineGen.forward + _f02sine (istftnet.py:146-186, 218-246), non-pulse branch.

    f0_frames: [B, T] frame-rate F0 (Generator.forward nearest-upsamples it by ``spf``,
    istftnet.py:351).  noise: [B, T*spf, 9] standard normal (the ``randn_like`` draw at
    :241).  ``rand_ini`` (:155-158) is not modelled: it only touches sample 0, and the
    linear down-sampling at :174 reads samples 300q+149 / 300q+150 only, so it never
    reaches the output.
    Returns sine_waves [B, T*spf, 9] and uv [B, T*spf, 1].
    """
    B, T = f0_frames.shape
    f0f = f0_frames.to(torch.float32)
    harm = torch.arange(1, N_HARM + 1, dtype=torch.float32, device=f0f.device)
    fn = f0f[:, :, None] * harm  # :227, per frame (nearest upsample keeps frame values)
    rad = (fn / SAMPLING_RATE) % 1  # :152
    # :174 linear interpolation by 1/spf: 0.5*rad[..149] + 0.5*rad[..150] == rad (frame value)
    if dtype == torch.float64:
        phase = torch.cumsum(rad.double(), dim=1) * 2 * np.pi  # :183
        phase_t = (phase.transpose(1, 2) * spf)
        up = F.interpolate(phase_t, scale_factor=spf, mode="linear").transpose(1, 2)  # :184
        sines = torch.sin(up)
    else:
        phase = torch.cumsum(rad, dim=1) * 2 * np.pi
        phase_t = (phase.transpose(1, 2) * spf)
        up = F.interpolate(phase_t, scale_factor=spf, mode="linear").transpose(1, 2)
        sines = torch.sin(up)
    sine_waves = sines.to(dtype) * SINE_AMP  # :230
    f0_up = f0f.repeat_interleave(spf, dim=1)[:, :, None]
    uv = (f0_up > VOICED_THRESHOLD).to(dtype)  # :143
    noise_amp = uv * NOISE_STD + (1 - uv) * SINE_AMP / 3  # :240
    sine_waves = sine_waves * uv + noise_amp * noise.to(dtype)  # :241,245
    return sine_waves, uv


def source_module(f0_frames, noise, P: Params, spf: int):
    """SourceModuleHnNSF.forward, istftnet.py:282-296 -> har_source [B, T*spf]."""
    sine_waves, _uv = sine_gen(f0_frames, noise, spf, P.dtype)
    merged = torch.tanh(F.linear(sine_waves, P.get("generator.m_source.l_linear.weight"),
                                 P.get("generator.m_source.l_linear.bias")))
    return merged[:, :, 0]


# ------------------------------------------------------------------ STFT / iSTFT
def hann_periodic(n: int, dtype) -> torch.Tensor:
    """scipy.signal.get_window('hann', n, fftbins=True), istftnet.py:89."""
    k = torch.arange(n, dtype=torch.float64)
    return (0.5 - 0.5 * torch.cos(2 * math.pi * k / n)).to(torch.float32).to(dtype)


def stft_transform(x: torch.Tensor, n_fft: int, hop: int, library: bool = False):
    """TorchSTFT.transform, istftnet.py:91-97: torch.stft(center=True, reflect pad,
    onesided) then abs / angle.  x: [B, L] -> mag, phase [B, n_fft/2+1, L/hop+1]."""
    dtype = x.dtype
    w = hann_periodic(n_fft, dtype).to(x.device)
    if library:  # the reference's own call, istftnet.py:92-97
        X = torch.stft(x, n_fft, hop, n_fft, window=w, return_complex=True)
        return torch.abs(X), torch.angle(X)
    xp = F.pad(x[:, None, :], (n_fft // 2, n_fft // 2), mode="reflect")[:, 0]
    frames = xp.unfold(1, n_fft, hop)  # [B, F, n_fft]
    n = torch.arange(n_fft, dtype=torch.float64, device=x.device)
    k = torch.arange(n_fft // 2 + 1, dtype=torch.float64, device=x.device)
    ang = 2 * math.pi * k[:, None] * n[None, :] / n_fft
    cosb, sinb = torch.cos(ang).to(dtype), torch.sin(ang).to(dtype)
    fw = frames * w
    re = torch.einsum("bfn,kn->bkf", fw, cosb)
    im = -torch.einsum("bfn,kn->bkf", fw, sinb)
    # DC and Nyquist of a real FFT have Im == +0.0 exactly (pocketfft/cuFFT r2c), so
    # torch.angle gives 0 or +pi there, never -pi.  [probed against torch.stft]
    im[:, 0] = 0.0
    im[:, -1] = 0.0
    return torch.sqrt(re * re + im * im), torch.atan2(im, re)


def istft_inverse(mag: torch.Tensor, phase: torch.Tensor, n_fft: int, hop: int, library: bool = False):
    """TorchSTFT.inverse, istftnet.py:99-104: torch.istft(mag*exp(i*phase)), center=True.
    Per frame irfft (imag of DC/Nyquist ignored), window, overlap-add, divide by the
    summed squared window, trim n_fft/2 both sides -> [B, 1, hop*(F-1)]."""
    dtype = mag.dtype
    dev = mag.device
    B, K, Fr = mag.shape
    w = hann_periodic(n_fft, dtype).to(dev)
    if library:  # the reference's own call, istftnet.py:100-104
        y = torch.istft(mag.float() * torch.exp(phase.float() * 1j), n_fft, hop, n_fft, window=w.float())
        return y.unsqueeze(-2).to(dtype)
    re, im = mag * torch.cos(phase), mag * torch.sin(phase)
    n = torch.arange(n_fft, dtype=torch.float64, device=dev)
    k = torch.arange(K, dtype=torch.float64, device=dev)
    ang = 2 * math.pi * k[:, None] * n[None, :] / n_fft
    c = torch.full((K,), 2.0, dtype=torch.float64, device=dev)
    c[0] = 1.0
    c[-1] = 1.0
    cosb = (c[:, None] * torch.cos(ang) / n_fft).to(dtype)
    sinb = (c[:, None] * torch.sin(ang) / n_fft).to(dtype)
    fr = torch.einsum("bkf,kn->bfn", re, cosb) - torch.einsum("bkf,kn->bfn", im, sinb)
    fr = fr * w
    total = hop * (Fr - 1) + n_fft
    y = torch.zeros(B, total, dtype=dtype, device=dev)
    env = torch.zeros(total, dtype=dtype, device=dev)
    for f in range(Fr):  # overlap-add (small Fr in tests; vectorised below for long ones)
        if Fr > 4096:
            break
        y[:, f * hop:f * hop + n_fft] += fr[:, f]
        env[f * hop:f * hop + n_fft] += w * w
    if Fr > 4096:
        y = F.fold(fr.transpose(1, 2), (1, total), (1, n_fft), stride=(1, hop))[:, 0, 0]
        env = F.fold((w * w)[None, :, None].expand(1, n_fft, Fr), (1, total), (1, n_fft),
                     stride=(1, hop))[0, 0, 0]
    s = n_fft // 2
    return (y[:, s:total - s] / env[s:total - s])[:, None, :]


# -------------------------------------------------------------------- generators
def generator_istftnet(x, s, f0_frames, P: Params, cfg, noise=None, har_source=None,
                       har=None, taps: Optional[dict] = None):
    """Generator.forward, istftnet.py:349-379."""
    g = "generator"
    n_fft, hop = cfg["gen_istft_n_fft"], cfg["gen_istft_hop_size"]
    rates, ks = cfg["upsample_rates"], cfg["upsample_kernel_sizes"]
    spf = int(np.prod(rates)) * hop
    nk = len(cfg["resblock_kernel_sizes"])
    if har is None:
        if har_source is None:
            har_source = source_module(f0_frames, noise, P, spf)
        mag, ph = stft_transform(har_source.to(P.dtype), n_fft, hop, P.library)
        har = torch.cat([mag, ph], dim=1)  # :356
    har = har.to(P.dtype)
    if taps is not None:
        taps["har_source"], taps["har"] = har_source, har
    for i in range(len(rates)):
        x = F.leaky_relu(x, LRELU_SLOPE)
        if i + 1 < len(rates):
            stride_f0 = int(np.prod(rates[i + 1:]))
            xs = F.conv1d(har, P.weight(f"{g}.noise_convs.{i}"), P.bias(f"{g}.noise_convs.{i}"),
                          stride=stride_f0, padding=(stride_f0 + 1) // 2)  # :333-335
            rk = 7
        else:
            xs = F.conv1d(har, P.weight(f"{g}.noise_convs.{i}"), P.bias(f"{g}.noise_convs.{i}"))
            rk = 11
        xs = adain_resblock1(xs, s, P, f"{g}.noise_res.{i}", rk)
        x = F.conv_transpose1d(x, P.weight(f"{g}.ups.{i}"), P.bias(f"{g}.ups.{i}"),
                               stride=rates[i], padding=(ks[i] - rates[i]) // 2)  # :319-321
        if i == len(rates) - 1:
            x = F.pad(x, (1, 0), mode="reflect")  # :345,:365
        x = x + xs
        acc = None
        for j, k_r in enumerate(cfg["resblock_kernel_sizes"]):
            y = adain_resblock1(x, s, P, f"{g}.resblocks.{i * nk + j}", k_r,
                                cfg["resblock_dilation_sizes"][j])
            acc = y if acc is None else acc + y
        x = acc / nk
        if taps is not None:
            taps[f"stage{i}"] = x
    x = F.leaky_relu(x)  # default slope 0.01 (!), :375
    x = F.conv1d(x, P.weight(g + ".conv_post"), P.bias(g + ".conv_post"), padding=3)
    if taps is not None:
        taps["conv_post"] = x
    nb = n_fft // 2 + 1
    spec, phase = torch.exp(x[:, :nb]), torch.sin(x[:, nb:])  # :377-378
    return istft_inverse(spec, phase, n_fft, hop, P.library)


def generator_hifigan(x, s, f0_frames, P: Params, cfg, noise=None, har_source=None,
                      taps: Optional[dict] = None):
    """Generator.forward, hifigan.py:318-346."""
    g = "generator"
    rates, ks = cfg["upsample_rates"], cfg["upsample_kernel_sizes"]
    spf = int(np.prod(rates))
    nk = len(cfg["resblock_kernel_sizes"])
    if har_source is None:
        har_source = source_module(f0_frames, noise, P, spf)
    if taps is not None:
        taps["har_source"] = har_source
    hs = har_source.to(P.dtype)[:, None, :]
    for i, u in enumerate(rates):
        x = snake(x, P.get(f"{g}.alphas.{i}"))  # :328
        if i + 1 < len(rates):
            stride_f0 = int(np.prod(rates[i + 1:]))
            xs = F.conv1d(hs, P.weight(f"{g}.noise_convs.{i}"), P.bias(f"{g}.noise_convs.{i}"),
                          stride=stride_f0, padding=(stride_f0 + 1) // 2)
            rk = 7
        else:
            xs = F.conv1d(hs, P.weight(f"{g}.noise_convs.{i}"), P.bias(f"{g}.noise_convs.{i}"))
            rk = 11
        xs = adain_resblock1(xs, s, P, f"{g}.noise_res.{i}", rk)
        x = F.conv_transpose1d(x, P.weight(f"{g}.ups.{i}"), P.bias(f"{g}.ups.{i}"), stride=u,
                               padding=u // 2 + u % 2, output_padding=u % 2)  # :291-293
        x = x + xs
        acc = None
        for j, k_r in enumerate(cfg["resblock_kernel_sizes"]):
            y = adain_resblock1(x, s, P, f"{g}.resblocks.{i * nk + j}", k_r,
                                cfg["resblock_dilation_sizes"][j])
            acc = y if acc is None else acc + y
        x = acc / nk
        if taps is not None:
            taps[f"stage{i}"] = x
    x = snake(x, P.get(f"{g}.alphas.{len(rates)}"))  # :342
    x = F.conv1d(x, P.weight(g + ".conv_post"), P.bias(g + ".conv_post"), padding=3)
    return torch.tanh(x)


# ----------------------------------------------------------------------- decoder
@torch.no_grad()
def decoder_forward(sd: Dict[str, torch.Tensor], cfg, asr, F0_curve, N, s, noise=None,
                    har_source=None, har=None, dtype=torch.float32,
                    taps: Optional[dict] = None, library: bool = False) -> torch.Tensor:
    """Decoder.forward in eval mode, istftnet.py:509-527 / hifigan.py:456-474.

    asr [B,512,T/2], F0_curve [B,T], N [B,T], s [B,style].  Exactly one of ``noise``
    ([B, T*spf, 9] standard normal), ``har_source`` ([B, T*spf]) or ``har``
    ([B,22,60T+1], istftnet only) supplies the excitation.  Returns [B,1,T*spf]."""
    P = Params(sd, dtype, library)
    asr, F0_curve, N, s = (t.to(dtype) for t in (asr, F0_curve, N, s))
    F0 = F.conv1d(F0_curve[:, None], P.weight("F0_conv"), P.bias("F0_conv"), stride=2, padding=1)
    Nn = F.conv1d(N[:, None], P.weight("N_conv"), P.bias("N_conv"), stride=2, padding=1)
    x = torch.cat([asr, F0, Nn], dim=1)
    x = adain_resblk1d(x, s, P, "encode", False)
    if taps is not None:
        taps["encode"] = x
    asr_res = F.conv1d(asr, P.weight("asr_res.0"), P.bias("asr_res.0"))
    for i in range(4):
        x = torch.cat([x, asr_res, F0, Nn], dim=1)  # res stays True until the upsample block
        x = adain_resblk1d(x, s, P, f"decode.{i}", upsample=(i == 3))
        if taps is not None:
            taps[f"decode{i}"] = x
    if cfg["kind"] == "istftnet":
        return generator_istftnet(x, s, F0_curve, P, cfg, noise, har_source, har, taps)
    return generator_hifigan(x, s, F0_curve, P, cfg, noise, har_source, taps)


# ----------------------------------------------------------------------- metrics
def snr_db(ref: torch.Tensor, test: torch.Tensor) -> float:
    ref, test = ref.double().flatten(), test.double().flatten()
    return float(10 * torch.log10((ref ** 2).sum() / ((ref - test) ** 2).sum().clamp_min(1e-300)))


def log_mel(wav: torch.Tensor, sr=16000, n_fft=2048, win=1200, hop=300, n_mels=80):
    """The reference's mel front end (meldataset.py:35-43: torchaudio MelSpectrogram
    n_mels 80 / n_fft 2048 / win 1200 / hop 300, log(1e-5+mel), (x+4)/4) restated with
    torch.stft and an HTK mel filterbank (torchaudio's defaults: power 2, htk, norm None, and
    sample_rate 16000 -- the reference never passes its 24 kHz rate, so the filters span 0..8000 "Hz" over
    the 1025 bins).  Pinned on the reference's own statements by oracle/make_golden_logmel.py ->
    tests/golden/logmel.npz (round 1 used sr=24000 here: 2.0 off the reference's values)."""
    wav = wav.reshape(-1, wav.shape[-1]).double()
    w = torch.hann_window(win, periodic=True, dtype=torch.float64)
    X = torch.stft(wav, n_fft, hop, win, window=w, center=True, pad_mode="reflect",
                   return_complex=True)
    power = X.abs() ** 2
    nfreq = n_fft // 2 + 1
    freqs = torch.linspace(0, sr // 2, nfreq, dtype=torch.float64)
    hz2mel = lambda f: 2595.0 * math.log10(1.0 + f / 700.0)
    m_pts = torch.linspace(hz2mel(0.0), hz2mel(sr / 2), n_mels + 2, dtype=torch.float64)
    f_pts = 700.0 * (10 ** (m_pts / 2595.0) - 1.0)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts[None, :] - freqs[:, None]
    down = -slopes[:, :-2] / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    fb = torch.clamp(torch.minimum(down, up), min=0)  # [nfreq, n_mels]
    mel = torch.einsum("bft,fm->bmt", power, fb)
    return (torch.log(1e-5 + mel) + 4) / 4


def log_mel_l1(ref: torch.Tensor, test: torch.Tensor) -> float:
    return float((log_mel(ref) - log_mel(test)).abs().mean())

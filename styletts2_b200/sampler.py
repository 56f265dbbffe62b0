"""Style sampler on the B200: the step that produces the style vector ``s``
the decoder and the prosody predictor consume.

Mirrors the reference boundary: ``DiffusionSampler(model.diffusion.diffusion, sampler=ADPM2Sampler(),
sigma_schedule=KarrasSchedule(sigma_min=0.0001, sigma_max=3.0, rho=9.0), clamp=False)`` called as
``sampler(noise=..., embedding=bert_dur, embedding_scale=..., features=ref_s, num_steps=...)``
(/root/reference/Demo/Inference_LibriTTS.ipynb cells 14 and 16; Modules/diffusion/sampler.py:549-585).
The module holds the denoiser network under the reference's state-dict names: ``StyleTransformer1d``
(multispeaker, ``context_features`` > 0) or ``Transformer1d`` (Modules/diffusion/modules.py:40-426), as
``build_model`` wires them (models.py:1057-1089, Configs/config*.yml ``diffusion.transformer``).
Forward-only, CUDA sm_100 tensors only, no CPU fallback (same rules as the decoder modules).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _cabi
from ._cabi import lib, check, ptr
from .decoder import _DecoderBase

_PREFIXES = ("diffusion.net.", "unet.", "net.")


class StyleSampler(_DecoderBase):
    """``StyleSampler(...)(noise, embedding=..., features=..., num_steps=5, embedding_scale=1.0) -> s_pred``

    Constructor kwargs are those of ``StyleTransformer1d`` / ``Transformer1d`` plus the sampler's
    (``sigma_min``, ``sigma_max``, ``rho`` of KarrasSchedule, ``sampler_rho`` of ADPM2Sampler,
    ``sigma_data`` of KDiffusion, ``clamp`` and ``num_steps`` of DiffusionSampler)."""
    KIND = _cabi.KIND_SAMPLER

    def __init__(self, channels: int = 256, context_embedding_features: int = 768, context_features: Optional[int] = 256,
                 num_layers: int = 3, num_heads: int = 8, head_features: int = 64, multiplier: int = 2,
                 embedding_max_length: int = 512, sigma_min: float = 0.0001, sigma_max: float = 3.0, rho: float = 9.0,
                 sampler_rho: float = 1.0, sigma_data: float = 0.2, clamp: bool = False, num_steps: Optional[int] = None):
        super().__init__()
        cfg = _cabi.St2Config()
        cfg.kind = _cabi.KIND_SAMPLER
        cfg.dim_in, cfg.style_dim, cfg.dim_out = int(channels), int(context_features or 0), int(context_embedding_features)
        cfg.n_resblock_kernels = int(num_layers)
        for i, v in enumerate((num_heads, head_features, multiplier, embedding_max_length)):
            cfg.resblock_kernel_sizes[i] = int(v)
        self.channels, self.context_features = int(channels), int(context_features or 0)
        self.embedding_features, self.embedding_max_length = int(context_embedding_features), int(embedding_max_length)
        self.sigma_min, self.sigma_max, self.rho = float(sigma_min), float(sigma_max), float(rho)
        self.sampler_rho, self.sigma_data, self.clamp, self.num_steps = float(sampler_rho), float(sigma_data), bool(clamp), num_steps
        self._setup(cfg)
        self.precision = "bf16"

    @torch.no_grad()
    def _reset_parameters(self):
        """torch defaults: nn.Linear / Conv1d U(+-1/sqrt(fan_in)), LayerNorm (1, 0), randn Fourier weights and
        embedding table (modules.py:629-636, 654-659)"""
        super()._reset_parameters()
        for name, p in self.named_parameters():
            if name.endswith("weights") or name.startswith("fixed_embedding."):
                p.normal_()
            elif p.dim() == 1 and ".attention.norm" in name and ".fc." not in name:
                p.fill_(1.0 if name.endswith("weight") else 0.0)

    def load_diffusion_state_dict(self, sd, strict: bool = True):
        """load the denoiser out of a full ``model.diffusion`` (AudioDiffusionConditional) state dict, which holds
        it twice, as ``diffusion.net.*`` and ``unet.*`` (models.py:1088-1089)"""
        for pref in _PREFIXES:
            sub = {}
            for k, v in sd.items():
                kk = k[7:] if k.startswith("module.") else k
                if kk.startswith(pref):
                    sub[kk[len(pref):]] = v
            if sub:
                return self.load_state_dict(sub, strict=strict)
        return self.load_state_dict(sd, strict=strict)

    # ------------------------------------------------------------------ calls
    def _opts(self, prec: str, num_steps: int, embedding_scale: float, step_noise=None, seed=None, item_offset: int = 0):
        o = _cabi.St2SamplerOpts()
        o.precision = {"fp32": _cabi.PREC_FP32, "bf16": _cabi.PREC_BF16}[prec]
        o.num_steps = int(num_steps)
        o.sigma_min, o.sigma_max, o.rho = self.sigma_min, self.sigma_max, self.rho
        o.sampler_rho, o.sigma_data, o.embedding_scale = self.sampler_rho, self.sigma_data, float(embedding_scale)
        o.clamp = int(self.clamp)
        o.step_noise = None if step_noise is None else step_noise.data_ptr()
        o.seed = int(self.seed if seed is None else seed) & (2 ** 64 - 1)
        o.item_offset = int(item_offset)
        return o

    def _check(self, x, embedding, features):
        if self.training:
            raise RuntimeError("styletts2_b200.StyleSampler is forward-only: call .eval() first")
        if not x.is_cuda:
            raise RuntimeError("styletts2_b200.StyleSampler needs CUDA tensors (sm_100a); there is no CPU path")
        lead = x.shape
        if x.dim() == 3 and x.shape[1] == 1:
            x = x[:, 0]
        if x.dim() != 2 or x.shape[1] != self.channels:
            raise RuntimeError(f"expected noise [B, 1, {self.channels}] or [B, {self.channels}], got {tuple(lead)}")
        if embedding is None or embedding.dim() != 3 or embedding.shape[0] != x.shape[0] or \
                embedding.shape[2] != self.embedding_features:
            raise RuntimeError(f"expected embedding [B, n_tokens, {self.embedding_features}]")
        if embedding.shape[1] > self.embedding_max_length:
            raise RuntimeError(f"n_tokens {embedding.shape[1]} exceeds embedding_max_length {self.embedding_max_length}")
        if self.context_features:
            if features is None or tuple(features.shape) != (x.shape[0], self.context_features):
                raise RuntimeError(f"context_features exists but no features [B, {self.context_features}] provided")
        elif features is not None:
            raise RuntimeError("Transformer1d (context_features=None) takes no features")
        dev = x.device
        f32 = lambda t: None if t is None else t.to(dev, torch.float32).contiguous()
        return f32(x), f32(embedding), f32(features), lead

    def _run(self, single, x, sigma, embedding, features, opts, prec):
        B, n = x.shape[0], embedding.shape[1]
        dev = x.device
        with self._lock, torch.cuda.device(dev):
            h = self._handle_for(dev)
            need = lib.st2_sampler_workspace_bytes(h, B, n, opts.precision, int(opts.embedding_scale != 1.0))
            if need < 0:
                check(int(need), "st2_sampler_workspace_bytes")
            ws = torch.empty(int(need), dtype=torch.uint8, device=dev)
            out = torch.empty(B, self.channels, dtype=torch.float32, device=dev)
            stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            if single:
                check(lib.st2_sampler_denoise(h, ptr(x), float(sigma), ptr(embedding), ptr(features), ptr(out), ptr(ws),
                                              int(need), B, n, C.byref(opts), stream), "st2_sampler_denoise")
            else:
                check(lib.st2_sampler_sample(h, ptr(x), ptr(embedding), ptr(features), ptr(out), ptr(ws), int(need), B, n,
                                             C.byref(opts), stream), "st2_sampler_sample")
            self.last_launches = int(lib.st2_decoder_last_launches(h))
        return out

    def forward(self, noise: torch.Tensor, num_steps: Optional[int] = None, embedding: Optional[torch.Tensor] = None,
                embedding_scale: float = 1.0, features: Optional[torch.Tensor] = None,
                step_noise: Optional[torch.Tensor] = None, precision: Optional[str] = None, seed: Optional[int] = None,
                item_offset: int = 0) -> torch.Tensor:
        """DiffusionSampler.forward (sampler.py:573-585).  ``step_noise`` [num_steps - 1, B, (1,) channels] injects the
        ``randn_like`` draws of the ADPM2 steps (sampler.py:510); without it they come from an in-kernel Philox
        stream keyed by (seed, item_offset + item, channel, step).  Returns the shape of ``noise``."""
        num_steps = self.num_steps if num_steps is None else num_steps
        if num_steps is None:
            raise RuntimeError("Parameter `num_steps` must be provided")
        x, emb, feat, lead = self._check(noise, embedding, features)
        sn = None
        if step_noise is not None:
            sn = step_noise.to(x.device, torch.float32).reshape(-1, x.shape[0], self.channels).contiguous()
            if sn.shape[0] < num_steps - 1:
                raise RuntimeError("step_noise needs num_steps - 1 draws")
        prec = precision or self.precision
        out = self._run(False, x, 0.0, emb, feat, self._opts(prec, num_steps, embedding_scale, sn, seed, item_offset), prec)
        if sn is None and seed is None:
            self.seed += 1   # fresh step noise on every call, like the reference's randn_like (sampler.py:510)
        return out.reshape(lead)

    def denoise_fn(self, x_noisy: torch.Tensor, sigma: float, embedding: Optional[torch.Tensor] = None,
                   embedding_scale: float = 1.0, features: Optional[torch.Tensor] = None,
                   precision: Optional[str] = None) -> torch.Tensor:
        """KDiffusion.denoise_fn(x_noisy, sigma=sigma, ...) (sampler.py:193-208): one denoiser evaluation"""
        x, emb, feat, lead = self._check(x_noisy, embedding, features)
        prec = precision or self.precision
        return self._run(True, x, sigma, emb, feat, self._opts(prec, 2, embedding_scale), prec).reshape(lead)

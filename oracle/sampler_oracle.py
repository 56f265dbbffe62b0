"""CPU restatement of the style sampler: ``DiffusionSampler`` with
``ADPM2Sampler`` and ``KarrasSchedule`` over the ``KDiffusion`` denoiser wrapping ``StyleTransformer1d``
(multispeaker) or ``Transformer1d``.

TEST INFRASTRUCTURE ONLY (imported by tests/, smoke() and bench.py's cpu_baseline leg; never by the
product path).  Pinned against the real reference by oracle/make_golden_sampler.py ->
tests/golden/sampler_*.npz (tests/test_oracle_golden.py).  References are to /root/reference.
"""
from __future__ import annotations

from math import pi, sqrt
from typing import Dict, Optional

import torch
import torch.nn.functional as F


def karras_sigmas(num_steps: int, sigma_min: float, sigma_max: float, rho: float) -> torch.Tensor:
    """KarrasSchedule.forward (Modules/diffusion/sampler.py:318-336): num_steps values + a trailing 0"""
    rho_inv = 1.0 / rho
    steps = torch.arange(num_steps, dtype=torch.float32)
    sig = (sigma_max ** rho_inv + (steps / (num_steps - 1)) * (sigma_min ** rho_inv - sigma_max ** rho_inv)) ** rho
    return F.pad(sig, pad=(0, 1), value=0.0)


def adpm2_sigmas(sigma, sigma_next, rho: float = 1.0):
    """ADPM2Sampler.get_sigmas (sampler.py:489-494); math.sqrt of 0-dim fp32 tensors, like the reference"""
    sigma_up = sqrt(sigma_next ** 2 * (sigma ** 2 - sigma_next ** 2) / sigma ** 2)
    sigma_down = sqrt(sigma_next ** 2 - sigma_up ** 2)
    sigma_mid = ((sigma ** (1 / rho) + sigma_down ** (1 / rho)) / 2) ** rho
    return sigma_up, sigma_down, sigma_mid


def ada_layer_norm(x, s, w, b, eps=1e-5):
    """AdaLayerNorm.forward (Modules/diffusion/modules.py:18-38) on x [B, n, F], s [B, S]"""
    h = F.linear(s, w, b)
    gamma, beta = h.chunk(2, dim=1)
    xh = F.layer_norm(x, (x.shape[-1],), eps=eps)
    return (1 + gamma[:, None, :]) * xh + beta[:, None, :]


class Net:
    """StyleTransformer1d / Transformer1d forward (modules.py:121-186 / 357-426)"""

    def __init__(self, sd: Dict[str, torch.Tensor], cfg):
        self.sd, self.cfg = sd, cfg
        self.style = bool(cfg["context_features"])

    def mapping(self, time, features):
        sd = self.sd
        # to_time: LearnedPositionalEmbedding + Linear + GELU (modules.py:629-652)
        t = time[:, None]
        freqs = t * sd["to_time.0.0.weights"][None, :] * 2 * pi
        four = torch.cat((t, freqs.sin(), freqs.cos()), dim=-1)
        m = F.gelu(F.linear(four, sd["to_time.0.1.weight"], sd["to_time.0.1.bias"]))
        if self.style:  # to_features: Linear + GELU (modules.py:108-114), summed (:139)
            m = m + F.gelu(F.linear(features, sd["to_features.0.weight"], sd["to_features.0.bias"]))
        m = F.gelu(F.linear(m, sd["to_mapping.0.weight"], sd["to_mapping.0.bias"]))
        return F.gelu(F.linear(m, sd["to_mapping.2.weight"], sd["to_mapping.2.bias"]))

    def norm(self, x, s, p):
        sd = self.sd
        if self.style:
            return ada_layer_norm(x, s, sd[p + ".fc.weight"], sd[p + ".fc.bias"])
        return F.layer_norm(x, (x.shape[-1],), sd[p + ".weight"], sd[p + ".bias"])

    def block(self, x, s, i):
        sd, H = self.sd, self.cfg["num_heads"]
        p = f"blocks.{i}.attention"
        # StyleAttention.forward (modules.py:268-278): separate norms for q and for k/v, same input
        q = F.linear(self.norm(x, s, p + ".norm"), sd[p + ".to_q.weight"])
        k, v = F.linear(self.norm(x, s, p + ".norm_context"), sd[p + ".to_kv.weight"]).chunk(2, dim=-1)
        B, n, _ = q.shape
        q, k, v = (t.reshape(B, n, H, -1).transpose(1, 2) for t in (q, k, v))
        # AttentionBase.forward (modules.py:512-526), no relative position bias (use_rel_pos False)
        sim = torch.einsum("bhnd,bhmd->bhnm", q, k) * (q.shape[-1] ** -0.5)
        o = torch.einsum("bhnm,bhmd->bhnd", sim.softmax(dim=-1), v).transpose(1, 2).reshape(B, n, -1)
        x = F.linear(o, sd[p + ".attention.to_out.weight"], sd[p + ".attention.to_out.bias"]) + x
        p = f"blocks.{i}.feed_forward"
        hmid = F.gelu(F.linear(x, sd[p + ".0.weight"], sd[p + ".0.bias"]))
        return F.linear(hmid, sd[p + ".2.weight"], sd[p + ".2.bias"]) + x

    def run(self, x, time, embedding, features):
        """StyleTransformer1d.run (modules.py:146-161): x [B, 1, C] -> [B, 1, C]"""
        m = self.mapping(time, features)[:, None, :]
        x = torch.cat([x.expand(-1, embedding.size(1), -1), embedding], dim=-1)
        for i in range(self.cfg["num_layers"]):
            x = self.block(x + m, features, i)
        x = x.mean(dim=1)
        return F.linear(x, self.sd["to_out.1.weight"][:, :, 0], self.sd["to_out.1.bias"])[:, None, :]

    def __call__(self, x, time, embedding, features, embedding_scale=1.0):
        """forward (modules.py:163-186), embedding_mask_proba = 0 (inference)"""
        if embedding_scale != 1.0:
            fixed = self.sd["fixed_embedding.embedding.weight"][: embedding.shape[1]][None].expand_as(embedding)
            out = self.run(x, time, embedding, features)
            out_masked = self.run(x, time, fixed, features)
            return out_masked + (out - out_masked) * embedding_scale
        return self.run(x, time, embedding, features)


def denoise(net: Net, x_noisy, sigma, sigma_data, **kw):
    """KDiffusion.denoise_fn (sampler.py:184-208), dynamic_threshold 0"""
    sigmas = torch.full((x_noisy.shape[0],), float(sigma), dtype=torch.float32)
    c_noise = torch.log(sigmas) * 0.25
    s3 = sigmas[:, None, None]
    c_skip = (sigma_data ** 2) / (s3 ** 2 + sigma_data ** 2)
    c_out = s3 * sigma_data * (sigma_data ** 2 + s3 ** 2) ** -0.5
    c_in = (s3 ** 2 + sigma_data ** 2) ** -0.5
    return c_skip * x_noisy + c_out * net(c_in * x_noisy, c_noise, **kw)


def sample(sd, cfg, noise, embedding, features=None, num_steps=5, embedding_scale=1.0, step_noise=None,
           sigma_min=1e-4, sigma_max=3.0, rho=9.0, sampler_rho=1.0, sigma_data=0.2, clamp=False,
           trace: Optional[list] = None):
    """DiffusionSampler.forward (sampler.py:573-585) with ADPM2Sampler.forward / .step (:496-519).
    step_noise [num_steps - 1, B, 1, C] replaces the randn_like draw of each step (:510)."""
    net = Net(sd, cfg)
    kw = dict(embedding=embedding, features=features, embedding_scale=embedding_scale)
    sigmas = karras_sigmas(num_steps, sigma_min, sigma_max, rho)
    x = sigmas[0] * noise
    for i in range(num_steps - 1):
        sigma, sigma_next = sigmas[i], sigmas[i + 1]
        sigma_up, sigma_down, sigma_mid = adpm2_sigmas(sigma, sigma_next, sampler_rho)
        d = (x - denoise(net, x, sigma, sigma_data, **kw)) / sigma
        x_mid = x + d * (sigma_mid - sigma)
        d_mid = (x_mid - denoise(net, x_mid, sigma_mid, sigma_data, **kw)) / sigma_mid
        x = x + d_mid * (sigma_down - sigma)
        x = x + (step_noise[i] if step_noise is not None else torch.randn_like(x)) * sigma_up
        if trace is not None:
            trace.append(x.clone())
    return x.clamp(-1.0, 1.0) if clamp else x

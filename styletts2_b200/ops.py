"""torch custom-op shell around the C ABI: ``torch.ops.styletts2_b200.decoder_forward``.

The op takes the id of a live Decoder module (its native handle is per device) plus the four
reference inputs and the optional injected excitation, and returns the waveform.  It exists so the
decoder shows up as ONE opaque operator to torch tooling (profilers, export); the arithmetic is all
inside libst2dec.so.
"""
from __future__ import annotations

import weakref
from typing import Optional

import torch

_MODULES = weakref.WeakValueDictionary()


def register_module(mod) -> int:
    _MODULES[mod._id] = mod
    return mod._id


@torch.library.custom_op("styletts2_b200::decoder_forward", mutates_args=())
def decoder_forward(asr: torch.Tensor, F0_curve: torch.Tensor, N: torch.Tensor, s: torch.Tensor,
                    noise: Optional[torch.Tensor], har_source: Optional[torch.Tensor],
                    har: Optional[torch.Tensor], module_id: int, precision: str,
                    item_offset: int = 0) -> torch.Tensor:
    mod = _MODULES.get(module_id)
    if mod is None:
        raise RuntimeError(f"styletts2_b200: no live Decoder with id {module_id}")
    return mod._forward_native(asr, F0_curve, N, s, noise=noise, har_source=har_source, har=har,
                               precision=precision, item_offset=item_offset)


@decoder_forward.register_fake
def _(asr, F0_curve, N, s, noise, har_source, har, module_id, precision, item_offset=0):
    mod = _MODULES.get(module_id)
    spf = mod.samples_per_frame if mod is not None else 300
    return asr.new_empty((asr.shape[0], 1, F0_curve.shape[1] * spf), dtype=torch.float32)

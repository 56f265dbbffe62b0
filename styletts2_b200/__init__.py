"""styletts2_b200 — B200-native (sm_100a) StyleTTS2 waveform decoder.

Only the decoder hot path of the reference is implemented:
``styletts2_b200.istftnet.Decoder`` and ``styletts2_b200.hifigan.Decoder`` are drop-ins for
``Modules.istftnet.Decoder`` / ``Modules.hifigan.Decoder`` as wired by ``models.py:build_model``.
All compute runs in hand-written CUDA kernels behind the C ABI of ``include/st2_decoder.h``;
importing this package without the built library raises.
"""
from . import _cabi  # noqa: F401  (raises ImportError when libst2dec.so is missing)
from .decoder import IstftnetDecoder, HifiganDecoder  # noqa: F401
from .predictor import F0NPredictor  # noqa: F401  (ProsodyPredictor.F0Ntrain)
from .sampler import StyleSampler  # noqa: F401  (ADPM2 style sampler over StyleTransformer1d)
from . import ops  # noqa: F401  (registers torch.ops.styletts2_b200.decoder_forward)

__all__ = ["IstftnetDecoder", "HifiganDecoder", "F0NPredictor", "StyleSampler", "build_decoder"]


def build_decoder(decoder_args, hidden_dim=512, style_dim=128, n_mels=80):
    """What ``models.py:1002-1032`` does for ``args.decoder``: pick the class from
    ``decoder.type`` and pass the YAML's decoder block.  ``decoder_args`` is a mapping / Munch."""
    get = (lambda k: decoder_args[k]) if isinstance(decoder_args, dict) else (lambda k: getattr(decoder_args, k))
    kind = get("type")
    common = dict(dim_in=hidden_dim, style_dim=style_dim, dim_out=n_mels,
                  resblock_kernel_sizes=get("resblock_kernel_sizes"),
                  upsample_rates=get("upsample_rates"),
                  upsample_initial_channel=get("upsample_initial_channel"),
                  resblock_dilation_sizes=get("resblock_dilation_sizes"),
                  upsample_kernel_sizes=get("upsample_kernel_sizes"))
    assert kind in ("istftnet", "hifigan"), "Decoder type unknown"  # models.py:1003
    if kind == "istftnet":
        return IstftnetDecoder(gen_istft_n_fft=get("gen_istft_n_fft"),
                               gen_istft_hop_size=get("gen_istft_hop_size"), **common)
    return HifiganDecoder(**common)

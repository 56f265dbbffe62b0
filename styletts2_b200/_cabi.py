"""ctypes binding of libst2dec.so (include/st2_decoder.h).

The library is built in-tree by ``styletts2_b200/build.py`` (``make`` + nvcc, sm_100a).  There
is no Python / CPU fallback: if the shared object is missing, importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libst2dec.so")

ST2_OK = 0
KIND_ISTFTNET, KIND_HIFIGAN, KIND_F0N, KIND_SAMPLER = 0, 1, 2, 3
PREC_FP32, PREC_BF16, PREC_FP32_TC = 0, 1, 2
DT_F32, DT_BF16 = 0, 1
ACT_NONE, ACT_LRELU, ACT_SNAKE = 0, 1, 2
MAX_UPS, MAX_RESK = 4, 4


class St2Config(C.Structure):
    _fields_ = [
        ("kind", C.c_int), ("dim_in", C.c_int), ("style_dim", C.c_int), ("dim_out", C.c_int),
        ("n_resblock_kernels", C.c_int), ("resblock_kernel_sizes", C.c_int * MAX_RESK),
        ("resblock_dilations", (C.c_int * 3) * MAX_RESK),
        ("n_ups", C.c_int), ("upsample_rates", C.c_int * MAX_UPS),
        ("upsample_kernel_sizes", C.c_int * MAX_UPS), ("upsample_initial_channel", C.c_int),
        ("gen_istft_n_fft", C.c_int), ("gen_istft_hop_size", C.c_int),
    ]


class St2Tensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data", C.c_void_p), ("ndim", C.c_int), ("shape", C.c_int64 * 4)]


class St2ForwardOpts(C.Structure):
    _fields_ = [
        ("precision", C.c_int), ("noise", C.c_void_p), ("har_source", C.c_void_p), ("har", C.c_void_p),
        ("seed", C.c_uint64), ("har_source_out", C.c_void_p), ("har_out", C.c_void_p),
        ("tap_decode", C.c_void_p), ("tap_stage", C.c_void_p * MAX_UPS),
        ("item_offset", C.c_int64), ("seed_dev", C.c_void_p),
        ("align_idx", C.c_void_p), ("align_len", C.c_int),
    ]


class St2SamplerOpts(C.Structure):
    _fields_ = [
        ("precision", C.c_int), ("num_steps", C.c_int),
        ("sigma_min", C.c_float), ("sigma_max", C.c_float), ("rho", C.c_float),
        ("sampler_rho", C.c_float), ("sigma_data", C.c_float), ("embedding_scale", C.c_float),
        ("clamp", C.c_int), ("step_noise", C.c_void_p), ("seed", C.c_uint64), ("item_offset", C.c_int64),
    ]


class St2ConvParams(C.Structure):
    _fields_ = [
        ("B", C.c_int),
        ("x", C.c_void_p), ("x_dtype", C.c_int), ("x_bstride", C.c_int64), ("ldx", C.c_int),
        ("Lin", C.c_int), ("cin", C.c_int),
        ("ss", C.c_void_p), ("ss_stride", C.c_int), ("act", C.c_int), ("slope", C.c_float),
        ("alpha", C.c_void_p),
        ("k", C.c_int), ("dil", C.c_int), ("pad", C.c_int), ("stride", C.c_int),
        ("w", C.c_void_p), ("bias", C.c_void_p), ("n_cols", C.c_int),
        ("y", C.c_void_p), ("y_dtype", C.c_int), ("y_bstride", C.c_int64), ("ldy", C.c_int),
        ("Lout", C.c_int),
        ("res", C.c_void_p), ("res_bstride", C.c_int64), ("ldres", C.c_int), ("res_shift", C.c_int),
        ("out_scale", C.c_float), ("accumulate", C.c_int),
        ("stats", C.c_void_p), ("n_tiles", C.c_int), ("algo_flops", C.c_double), ("bias_bstride", C.c_int64),
        ("n_tile", C.c_int),
    ]


# every symbol include/st2_decoder.h declares: (restype, argtypes)
SYMBOLS = {
    "st2_decoder_create": (C.c_int, [C.POINTER(St2Config), C.POINTER(C.c_void_p)]),
    "st2_decoder_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int]),
    "st2_decoder_load": (C.c_int, [C.c_void_p, C.POINTER(St2Tensor), C.c_int]),
    "st2_decoder_num_keys": (C.c_int, [C.c_void_p]),
    "st2_decoder_key": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                  C.POINTER(C.c_int64)]),
    "st2_decoder_samples_per_frame": (C.c_int, [C.c_void_p]),
    "st2_decoder_workspace_bytes": (C.c_int64, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "st2_decoder_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int,
                                      C.POINTER(St2ForwardOpts), C.c_void_p]),
    "st2_f0n_workspace_bytes": (C.c_int64, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "st2_f0n_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "st2_sampler_workspace_bytes": (C.c_int64, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]),
    "st2_sampler_sample": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                     C.c_int, C.c_int, C.POINTER(St2SamplerOpts), C.c_void_p]),
    "st2_sampler_denoise": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_int64, C.c_int, C.c_int, C.POINTER(St2SamplerOpts), C.c_void_p]),
    "st2_wave_to_pcm16": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]),
    "st2_alignment_indices": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "st2_expand_alignment": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "st2_decoder_last_launches": (C.c_int64, [C.c_void_p]),
    "st2_decoder_destroy": (None, [C.c_void_p]),
    "st2_last_error": (C.c_char_p, []),
    "st2_version": (C.c_char_p, []),
    "st2_conv_simt": (C.c_int, [C.POINTER(St2ConvParams), C.c_void_p]),
    "st2_conv_umma": (C.c_int, [C.POINTER(St2ConvParams), C.c_void_p]),
    "st2_conv_simt_dense_launches": (C.c_int64, [C.c_int]),
    "st2_conv_umma_multi": (C.c_int, [C.POINTER(St2ConvParams), C.c_int, C.c_void_p]),
    "st2_umma_packed_elems": (C.c_int64, [C.c_int, C.c_int, C.c_int]),
    "st2_umma_pack_weights": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "st2_umma_pack_weights_nt": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "st2_debug_pack_conv": (C.c_int, [C.POINTER(St2Tensor), C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(C.c_int),
                                      C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "st2_prof_enable": (C.c_int, [C.c_int]),
    "st2_prof_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_double), C.c_int]),
    "st2_umma_wait_profile": (C.c_int, [C.POINTER(C.c_uint64), C.c_int]),
    "st2_umma_reload_env": (C.c_int, []),
    "st2_log_mel": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]),
    "st2_log_mel_frames": (C.c_int, [C.c_int64]),
    "st2_conv_tile_rows": (C.c_int, [C.c_int]),
    "st2_source_har": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_float, C.c_int,
                                 C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "st2_source_har_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_float, C.c_int,
                                        C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "st2_stft_har": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "st2_stft_har16": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "st2_istft_head": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
}


class St2Error(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python styletts2_b200/build.py` "
            "(nvcc, sm_100a).  styletts2_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc: int, what: str = ""):
    if rc != ST2_OK:
        msg = lib.st2_last_error().decode(errors="replace")
        raise St2Error(f"{what or 'libst2dec'} failed (code {rc}): {msg}")
    return rc


def ptr(t):
    """data pointer of a torch tensor (or None) as c_void_p."""
    return None if t is None else C.c_void_p(t.data_ptr())

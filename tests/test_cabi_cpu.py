"""CPU: the C-ABI library loads, exports every symbol include/st2_decoder.h declares, and its
host-side logic (config validation, key list, weight-norm folding, polyphase expansion, UMMA
blob packing) is right.  No compute entry point is called: there is no GPU here."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from styletts2_b200 import _cabi
from styletts2_b200._cabi import lib, check

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "st2_decoder.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(st2_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 20
    raw = C.CDLL(_cabi.LIB_PATH)
    for n in sorted(names):
        assert hasattr(raw, n), f"{n} declared in st2_decoder.h but not exported"
    assert names == set(_cabi.SYMBOLS), names ^ set(_cabi.SYMBOLS)
    assert b"sm_100a" in lib.st2_version()


def test_struct_layouts_match_header_sizes():
    # St2ConvParams: ints/pointers/int64 interleaved; ctypes applies the same natural alignment as C
    assert C.sizeof(_cabi.St2ConvParams) % 8 == 0
    assert C.sizeof(_cabi.St2Tensor) == 8 + 8 + 8 + 32
    assert C.sizeof(_cabi.St2ForwardOpts) == 56 + 8 + 32 + 8 + 8 + 8 + 8  # + item_offset, seed_dev, align_idx, align_len (+ padding)


@pytest.mark.parametrize("kind", ["istftnet", "hifigan"])
def test_key_list_matches_reference(kind):
    import styletts2_b200 as st2
    dec = (st2.IstftnetDecoder if kind == "istftnet" else st2.HifiganDecoder)(style_dim=128)
    ref = json.load(open(os.path.join(ROOT, "tests", "golden", f"keys_{kind}.json")))
    assert {k: list(v.shape) for k, v in dec.state_dict().items()} == ref
    assert dec.samples_per_frame == 300


@pytest.mark.parametrize("kind", ["istftnet", "hifigan"])
def test_module_default_init_matches_reference_constructor(kind):
    """the drop-in module's own random init (what a user gets from ``Decoder(...)`` before loading a
    checkpoint) has the per-layer statistics of the reference's constructor (tests/golden/init_stats_*.json):
    kaiming-uniform ``weight_v`` everywhere, because ``init_weights`` never lands on weight-normed layers"""
    import styletts2_b200 as st2
    import synthetic as synth
    torch.manual_seed(0)
    dec = (st2.IstftnetDecoder if kind == "istftnet" else st2.HifiganDecoder)(style_dim=128)
    ref = json.load(open(os.path.join(ROOT, "tests", "golden", f"init_stats_{kind}.json")))
    sd = dec.state_dict()
    mine = synth.init_stats(sd)
    for k in ("generator.resblocks.3.convs1.0.weight_v", "generator.ups.0.weight_v", "generator.conv_post.weight_v",
              "decode.0.conv1.weight_v", "generator.noise_convs.0.weight", "encode.norm1.fc.weight"):
        assert abs(mine[k][0] / ref[k][0] - 1) < 0.05, (k, mine[k], ref[k])
        assert abs(mine[k][1] / ref[k][1] - 1) < 0.05, (k, mine[k], ref[k])
    for k, v in mine.items():
        if k.endswith("weight_g"):
            assert v[0] < 1e-6 and abs(v[2] - 1) < 1e-6, k


@pytest.mark.parametrize("variant", ["style", "lj"])
def test_sampler_key_list_matches_reference(variant):
    import styletts2_b200 as st2
    import synthetic as synth
    cfg = synth.SAMPLER_CFG if variant == "style" else synth.SAMPLER_CFG_LJ
    m = st2.StyleSampler(**cfg)
    ref = json.load(open(os.path.join(ROOT, "tests", "golden", f"keys_sampler_{variant}.json")))
    assert {k: list(v.shape) for k, v in m.state_dict().items()} == ref
    with pytest.raises(_cabi.St2Error, match="head_features"):
        st2.StyleSampler(head_features=32)
    with pytest.raises(RuntimeError, match="CUDA"):  # no CPU path, fails loudly
        m(torch.zeros(1, 1, 256), num_steps=3, embedding=torch.zeros(1, 4, 768),
          features=torch.zeros(1, 256) if variant == "style" else None)
    assert C.sizeof(_cabi.St2SamplerOpts) == 64


def test_config_validation_and_errors():
    import styletts2_b200 as st2
    with pytest.raises(_cabi.St2Error, match="n_fft"):
        st2.IstftnetDecoder(style_dim=128, gen_istft_n_fft=16, gen_istft_hop_size=4)
    with pytest.raises(_cabi.St2Error):
        st2.IstftnetDecoder(style_dim=128, upsample_initial_channel=256)
    with pytest.raises(ValueError):
        st2.IstftnetDecoder(style_dim=128, upsample_rates=[10, 6], upsample_kernel_sizes=[20])
    dec = st2.IstftnetDecoder(style_dim=128)
    with pytest.raises(RuntimeError, match="CUDA"):  # no CPU path, fails loudly
        dec(torch.zeros(1, 512, 4), torch.zeros(1, 8), torch.zeros(1, 8), torch.zeros(1, 128))
    h = C.c_void_p()
    assert lib.st2_decoder_create(None, C.byref(h)) == -1
    assert b"null" in lib.st2_last_error()


def test_load_state_dict_is_strict_and_strips_module_prefix():
    import styletts2_b200 as st2
    import synthetic as synth
    dec = st2.IstftnetDecoder(**{k: v for k, v in synth.ISTFTNET_CFG.items() if k != "kind"})
    sd = synth.synth_state_dict(synth.ISTFTNET_CFG)
    dec.load_state_dict({"module." + k: v for k, v in sd.items()}, strict=True)
    assert torch.equal(dec.state_dict()["decode.3.pool.weight_v"], sd["decode.3.pool.weight_v"])
    bad = dict(sd)
    bad.pop("generator.conv_post.weight_g")
    with pytest.raises(RuntimeError, match="conv_post.weight_g"):
        dec.load_state_dict(bad, strict=True)


def _tensors(d):
    arr = (_cabi.St2Tensor * len(d))()
    keep = []
    for i, (k, v) in enumerate(d.items()):
        v = v.contiguous().float()
        kb = k.encode()
        keep += [v, kb]
        arr[i].name, arr[i].data, arr[i].ndim = kb, v.data_ptr(), v.dim()
        for j in range(v.dim()):
            arr[i].shape[j] = v.shape[j]
    return arr, keep


def _fold(g, v):
    return v * (g / v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, 1, 1))


def _pack(d, prefix, transposed, K, s, pd, cap):
    arr, keep = _tensors(d)
    w = torch.zeros(cap)
    b = torch.zeros(4096)
    k, p_, ci, nc = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    check(lib.st2_debug_pack_conv(arr, len(d), prefix.encode(), transposed, K, s, pd, w.data_ptr(), cap,
                                  b.data_ptr(), C.byref(k), C.byref(p_), C.byref(ci), C.byref(nc)))
    return w[: k.value * ci.value * nc.value].reshape(k.value, ci.value, nc.value), b[: nc.value], p_.value


@pytest.mark.parametrize("K,s,pd,op", [(20, 10, 5, 0), (12, 6, 3, 0), (10, 5, 3, 1), (6, 3, 2, 1), (4, 2, 1, 0)])
def test_polyphase_expansion_equals_conv_transpose(K, s, pd, op):
    """every (kernel, stride, padding) pair of both decoders (istftnet.py:319-321, hifigan.py:291-293)"""
    g_ = torch.Generator().manual_seed(K)
    cin, cout, L = 24, 16, 13
    v = torch.randn(cin, cout, K, generator=g_)
    g = torch.rand(cin, 1, 1, generator=g_) + 0.5
    b = torch.randn(cout, generator=g_)
    wk, bo, pad = _pack({"u.weight_v": v, "u.weight_g": g, "u.bias": b}, "u", 1, K, s, pd, 8 * cin * s * cout)
    x = torch.randn(2, cin, L, generator=g_)
    ref = F.conv_transpose1d(x, _fold(g, v), b, stride=s, padding=pd, output_padding=op)
    y = F.conv1d(x, wk.permute(2, 1, 0).contiguous(), bo, padding=pad)
    y = y.reshape(2, s, cout, L).permute(0, 2, 3, 1).reshape(2, cout, L * s)  # rows of s*Cout == s rows of Cout
    assert ref.shape == y.shape
    assert (ref - y).abs().max().item() < 2e-6


def test_weight_norm_fold_conv1d():
    g_ = torch.Generator().manual_seed(1)
    v = torch.randn(16, 24, 7, generator=g_)
    g = torch.rand(16, 1, 1, generator=g_) + 0.5
    b = torch.randn(16, generator=g_)
    wk, bo, pad = _pack({"c.weight_v": v, "c.weight_g": g, "c.bias": b}, "c", 0, 7, 1, 3, 7 * 24 * 16)
    assert (wk.permute(2, 1, 0) - _fold(g, v)).abs().max().item() < 1e-6
    assert torch.equal(bo, b) and pad == 3


def test_umma_blob_layout():
    """blob[nt][chunk][tap][v][n][e] = bf16(w[tap][chunk*64 + v*8 + e][nt*128 + n]), zero channel padding"""
    k, cin, cin_pad, n = 3, 70, 128, 256
    g_ = torch.Generator().manual_seed(2)
    w = torch.randn(k, cin, n, generator=g_)
    elems = lib.st2_umma_packed_elems(k, cin_pad, n)
    assert elems == (n // 128) * (cin_pad // 64) * k * 8 * 128 * 8
    out = torch.zeros(elems, dtype=torch.int16)
    check(lib.st2_umma_pack_weights(w.data_ptr(), k, cin, cin_pad, n, out.data_ptr()))
    blob = out.view(torch.bfloat16).float().reshape(n // 128, cin_pad // 64, k, 8, 128, 8)
    wp = torch.zeros(k, cin_pad, n)
    wp[:, :cin] = w.to(torch.bfloat16).float()
    expect = wp.reshape(k, cin_pad // 64, 8, 8, n // 128, 128).permute(4, 1, 0, 2, 5, 3)
    assert torch.equal(blob, expect)
    assert lib.st2_umma_pack_weights(w.data_ptr(), k, cin, 100, n, out.data_ptr()) == -1


def test_wav_framing_reads_back_with_the_wave_module():
    """output.wav_frame (row N4): the bytes are a valid mono 16-bit PCM .wav that the standard library reads back
    sample for sample"""
    import io
    import wave
    import numpy as np
    import torch
    from styletts2_b200.output import wav_frame
    pcm = torch.from_numpy((np.sin(np.arange(2401) * 0.01) * 30000).astype(np.int16))
    blob = wav_frame(pcm, 24000)
    assert len(blob) == 44 + 2 * pcm.numel() and blob[:4] == b"RIFF" and blob[8:12] == b"WAVE"
    with wave.open(io.BytesIO(blob)) as w:
        assert (w.getnchannels(), w.getsampwidth(), w.getframerate(), w.getnframes()) == (1, 2, 24000, 2401)
        back = np.frombuffer(w.readframes(2401), dtype="<i2")
    assert np.array_equal(back, pcm.numpy())
    assert wav_frame(pcm.numpy(), 16000)[24:28] == (16000).to_bytes(4, "little")
    import pytest
    with pytest.raises(ValueError):
        wav_frame(pcm.float())

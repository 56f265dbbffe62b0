"""Drop-in ``Decoder`` modules backed by libst2dec.so (hand-written sm_100a CUDA).

Mirrors the reference boundary:
  * constructor kwargs   /root/reference/Modules/istftnet.py:467-473, Modules/hifigan.py:416-421
  * ``forward(asr, F0_curve, N, s) -> [B, 1, 300*T]``   istftnet.py:498-527, hifigan.py:445-474
  * ``state_dict`` keys  old-style weight_norm ``*.weight_g`` / ``*.weight_v`` (istftnet.py:5)
so ``models.py:1006`` / ``:1021`` can import this class instead.  Weight-norm is folded when the
weights are (re)loaded into the native handle, not on every forward (the reference recomputes it
70 times per call).  Forward-only: ``.train()`` mode raises (the reference's training branch at
istftnet.py:499-507 hard-codes ``.to('cuda')`` smoothing filters and needs autograd).

No torch / CPU fallback exists: tensors must live on a CUDA device of compute capability 10.x.
"""
from __future__ import annotations

import ctypes as C
import math
import threading
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn as nn

from . import _cabi
from ._cabi import lib, check, ptr

_HANDLES: Dict[int, "_DecoderBase"] = {}
_NEXT_ID = [1]


def _mk_config(kind: int, dim_in, style_dim, dim_out, resblock_kernel_sizes, upsample_rates,
               upsample_initial_channel, resblock_dilation_sizes, upsample_kernel_sizes,
               gen_istft_n_fft=0, gen_istft_hop_size=0) -> _cabi.St2Config:
    cfg = _cabi.St2Config()
    cfg.kind = kind
    cfg.dim_in, cfg.style_dim, cfg.dim_out = int(dim_in), int(style_dim), int(dim_out)
    ks = list(resblock_kernel_sizes)
    if len(ks) > _cabi.MAX_RESK or len(upsample_rates) > _cabi.MAX_UPS:
        raise ValueError("too many resblock kernels / upsample stages")
    if len(upsample_rates) != len(upsample_kernel_sizes):
        raise ValueError("upsample_rates and upsample_kernel_sizes differ in length")
    cfg.n_resblock_kernels = len(ks)
    for i, k in enumerate(ks):
        cfg.resblock_kernel_sizes[i] = int(k)
        d = list(resblock_dilation_sizes[i])
        if len(d) != 3:
            raise ValueError("AdaINResBlock1 takes exactly three dilations")
        for j in range(3):
            cfg.resblock_dilations[i][j] = int(d[j])
    cfg.n_ups = len(upsample_rates)
    for i, (u, k) in enumerate(zip(upsample_rates, upsample_kernel_sizes)):
        cfg.upsample_rates[i] = int(u)
        cfg.upsample_kernel_sizes[i] = int(k)
    cfg.upsample_initial_channel = int(upsample_initial_channel)
    cfg.gen_istft_n_fft, cfg.gen_istft_hop_size = int(gen_istft_n_fft), int(gen_istft_hop_size)
    return cfg


class _Node(nn.Module):
    """anonymous container so that parameters get the reference's dotted names."""


class _DecoderBase(nn.Module):
    KIND = -1

    def _setup(self, cfg: _cabi.St2Config):
        self._cfg = cfg
        h = C.c_void_p()
        check(lib.st2_decoder_create(C.byref(cfg), C.byref(h)), "st2_decoder_create")
        try:
            n = lib.st2_decoder_num_keys(h)
            self._keys: List[str] = []
            name, ndim, shape = C.c_char_p(), C.c_int(), (C.c_int64 * 4)()
            for i in range(n):
                check(lib.st2_decoder_key(h, i, C.byref(name), C.byref(ndim), shape))
                key, shp = name.value.decode(), tuple(int(shape[d]) for d in range(ndim.value))
                self._keys.append(key)
                self._register(key, shp)
            self.samples_per_frame = lib.st2_decoder_samples_per_frame(h)
        finally:
            lib.st2_decoder_destroy(h)
        self._reset_parameters()
        self._native: Dict[int, C.c_void_p] = {}     # device index -> handle
        self._native_version: Dict[int, int] = {}
        self._id = _NEXT_ID[0]
        _NEXT_ID[0] += 1
        self.precision = "bf16"
        # "fp32tc" (the fp32 parity path with its dense convs on the tensor cores as 3 x bf16 products) needs extra weight
        # blobs: set fp32_tc = True BEFORE the first forward / after changing it the weights are re-packed
        self.fp32_tc = False
        self._native_tc: Dict[int, bool] = {}
        self._schedule: Dict[bytes, int] = {}   # set_schedule(): options re-applied to every new native handle
        self.seed = 0
        # a batch whose workspace would exceed this is decoded in micro-batches (utterances are
        # independent, SURVEY 8d: config C5 is 512 x 2400 frames); 0 = never split
        self.workspace_limit_bytes = 32 << 30
        # replay the forward as a CUDA graph (one per device / batch shape / precision, static buffers;
        # inputs are copied in, the waveform is copied out): removes ~150 launch gaps per forward
        self.cuda_graphs = False
        self.max_graphs = 4
        self._graphs: Dict[tuple, dict] = {}
        # one forward at a time per module (include/st2_decoder.h): the handle table, the graph cache, the seed and the
        # handle's launch count are shared by the threads of a one-thread-per-device caller (train_finetune.py:96-101);
        # launches are asynchronous, so devices still overlap
        self._lock = threading.RLock()
        self.eval()

    # ------------------------------------------------------------------ parameters
    def _register(self, key: str, shape):
        mod = self
        parts = key.split(".")
        for p in parts[:-1]:
            if p not in mod._modules:
                mod.add_module(p, _Node())
            mod = mod._modules[p]
        mod.register_parameter(parts[-1], nn.Parameter(torch.empty(shape), requires_grad=False))

    @torch.no_grad()
    def _reset_parameters(self):
        """Same distributions as the reference's constructors leave behind: torch's default
        kaiming-uniform U(+-1/sqrt(fan_in)) for every weight and bias, ``weight_v`` of the weight-normed
        layers included (``init_weights``, Modules/utils.py:1-4 via istftnet.py:38,48,343-344, writes the
        derived ``.weight`` attribute of an old-style weight-normed module, which the next forward
        recomputes from g and v, so its N(0, 0.01) never lands), weight_g = ||v||, Snake alpha = 1."""
        sd = dict(self.named_parameters())
        for key, p in sd.items():
            leaf = key.rsplit(".", 1)[-1]
            if leaf.startswith("alpha") or ".alphas." in key or ".alpha1." in key or ".alpha2." in key:
                p.fill_(1.0)
            elif leaf == "weight_v":
                fan_in = p.shape[1] * p.shape[2]
                p.uniform_(-1 / math.sqrt(fan_in), 1 / math.sqrt(fan_in))
            elif leaf == "weight":
                fan_in = p[0].numel()
                p.uniform_(-1 / math.sqrt(fan_in), 1 / math.sqrt(fan_in))
        for key, p in sd.items():
            leaf = key.rsplit(".", 1)[-1]
            if leaf == "weight_g":
                v = sd[key[:-1] + "v"]
                p.copy_(v.reshape(v.shape[0], -1).norm(dim=1).reshape(p.shape))
            elif leaf == "bias":
                base = key[: -len("bias")]
                w = sd.get(base + "weight_v", sd.get(base + "weight"))
                fan_in = w[0].numel() if w is not None and w.dim() > 1 else 1
                p.uniform_(-1 / math.sqrt(fan_in), 1 / math.sqrt(fan_in))

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        # tolerate the DDP "module." prefix like models.py:1126-1142 / the notebooks' k[7:]
        pref = prefix + "module."
        for k in [k for k in state_dict if k.startswith(pref)]:
            state_dict[prefix + k[len(pref):]] = state_dict.pop(k)
        return super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    def _weights_version(self) -> int:
        return sum(p._version for p in self.parameters())

    def _handle_for(self, device: torch.device) -> C.c_void_p:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        ver = self._weights_version()
        h = self._native.get(idx)
        want_tc = bool(getattr(self, "fp32_tc", False))
        if h is not None and self._native_version.get(idx) == ver and self._native_tc.get(idx, False) == want_tc:
            return h
        with torch.cuda.device(idx):
            if h is not None and self._native_tc.get(idx, False) != want_tc:
                lib.st2_decoder_destroy(h)      # the option changes what load packs: start from a fresh handle
                h = None
                for k in [k for k in self._graphs if k[0] == idx]:
                    del self._graphs[k]
            if h is None:
                h = C.c_void_p()
                check(lib.st2_decoder_create(C.byref(self._cfg), C.byref(h)), "st2_decoder_create")
                if want_tc:
                    check(lib.st2_decoder_set_option(h, b"fp32_tc", 1), "st2_decoder_set_option")
                for name, v in self._schedule.items():
                    check(lib.st2_decoder_set_option(h, name, v), "st2_decoder_set_option")
                self._native[idx] = h
                self._native_tc[idx] = want_tc
            sd = {k: v.detach().to("cpu", torch.float32).contiguous() for k, v in self.state_dict().items()}
            arr = (_cabi.St2Tensor * len(sd))()
            keep = []
            for i, (k, v) in enumerate(sd.items()):
                kb = k.encode()
                keep.append(kb)
                arr[i].name = kb
                arr[i].data = v.data_ptr()
                arr[i].ndim = v.dim()
                for d in range(v.dim()):
                    arr[i].shape[d] = v.shape[d]
            check(lib.st2_decoder_load(h, arr, len(sd)), "st2_decoder_load")
        self._native_version[idx] = ver
        return h

    def set_schedule(self, fork: Optional[int] = None, group: Optional[int] = None) -> None:
        """Schedule switches of the native forward (``st2_decoder_set_option``; results are bit-identical for every
        value).  ``fork``: 0 one stream, 1 (default) the excitation branch of a single-utterance-sized forward on a
        second stream, 2 for every size.  ``group``: 0 / 1 (default) independent same-shaped convs of such a forward as
        one launch.  Captured graphs are dropped: they hold the old schedule."""
        with self._lock:
            for name, v in ((b"fork", fork), (b"group", group)):
                if v is None:
                    continue
                self._schedule[name] = int(v)
                for h in self._native.values():
                    check(lib.st2_decoder_set_option(h, name, int(v)), "st2_decoder_set_option")
            self._graphs.clear()

    def __del__(self):
        try:
            for h in getattr(self, "_native", {}).values():
                lib.st2_decoder_destroy(h)
        except Exception:
            pass

    # --------------------------------------------------------------------- forward
    def forward(self, asr: torch.Tensor, F0_curve: torch.Tensor, N: torch.Tensor, s: torch.Tensor, *,
                noise: Optional[torch.Tensor] = None, har_source: Optional[torch.Tensor] = None,
                har: Optional[torch.Tensor] = None, precision: Optional[str] = None,
                return_excitation: bool = False, return_taps: bool = False, item_offset: int = 0,
                alignment: Optional[torch.Tensor] = None):
        """Reference signature plus keyword-only extras; dispatches through the registered custom
        op ``torch.ops.styletts2_b200.decoder_forward`` (see :meth:`_forward_native`).
        ``item_offset`` is the global index of row 0 when the call decodes a slice of a larger batch
        (the in-kernel noise is keyed by (seed, global item, sample)).  ``alignment`` (int32
        ``[B, T/2]`` from :func:`styletts2_b200.alignment.alignment_indices`) fuses the reference's
        ``asr = t_en @ pred_aln_trg``: the first argument is then the token-rate ``t_en``."""
        if return_excitation or return_taps or alignment is not None:
            return self._forward_native(asr, F0_curve, N, s, noise=noise, har_source=har_source, har=har,
                                        precision=precision, return_excitation=return_excitation,
                                        return_taps=return_taps, item_offset=item_offset, alignment=alignment)
        from . import ops
        return torch.ops.styletts2_b200.decoder_forward(asr, F0_curve, N, s, noise, har_source, har,
                                                        ops.register_module(self),
                                                        precision or self.precision, item_offset)

    def _forward_native(self, asr: torch.Tensor, F0_curve: torch.Tensor, N: torch.Tensor, s: torch.Tensor, *,
                        noise: Optional[torch.Tensor] = None, har_source: Optional[torch.Tensor] = None,
                        har: Optional[torch.Tensor] = None, precision: Optional[str] = None,
                        return_excitation: bool = False, return_taps: bool = False, item_offset: int = 0,
                        alignment: Optional[torch.Tensor] = None):
        """``Decoder.forward(asr, F0_curve, N, s)`` (istftnet.py:498).  Extra keyword-only inputs
        inject the excitation for reproducibility: ``noise`` [B, 300T, 9] (the randn_like draw of
        SineGen), ``har_source`` [B, 300T] (output of m_source) or ``har`` [B, 22, 60T+1]."""
        if self.training:
            raise RuntimeError("styletts2_b200.Decoder is forward-only: call .eval() first")
        if not asr.is_cuda:
            raise RuntimeError("styletts2_b200.Decoder needs CUDA tensors (sm_100a); there is no CPU path")
        with self._lock:
            return self._forward_locked(asr, F0_curve, N, s, noise=noise, har_source=har_source, har=har,
                                        precision=precision, return_excitation=return_excitation,
                                        return_taps=return_taps, item_offset=item_offset, alignment=alignment)

    def _forward_locked(self, asr, F0_curve, N, s, *, noise, har_source, har, precision, return_excitation,
                        return_taps, item_offset, alignment):
        if asr.dim() != 3 or F0_curve.dim() != 2 or N.dim() != 2 or s.dim() != 2:
            raise RuntimeError("expected asr [B,C,T/2], F0_curve [B,T], N [B,T], s [B,style]")
        B, Cin, T2 = asr.shape
        align_len = 0
        if alignment is not None:  # asr is t_en [B, C, Ltext]; frames come from the alignment
            if alignment.dim() != 2 or alignment.shape[0] != B or not alignment.is_cuda:
                raise RuntimeError("alignment must be a CUDA int32 tensor [B, T/2] (alignment_indices)")
            align_len, T2 = T2, alignment.shape[1]
            alignment = alignment.to(asr.device, torch.int32).contiguous()
        T = F0_curve.shape[1]
        if T != 2 * T2 or N.shape != F0_curve.shape or F0_curve.shape[0] != B or s.shape[0] != B:
            # the reference fails in torch.cat at istftnet.py:513 with a size error
            raise RuntimeError(f"Sizes of tensors must match: asr has {T2} frames, F0_curve/N have "
                               f"{T} and {N.shape[1]} (need 2x), batch {B}")
        if Cin != self._cfg.dim_in or s.shape[1] != self._cfg.style_dim:
            raise RuntimeError("channel / style dimension mismatch")
        prec = precision or self.precision
        prec_id = {"fp32": _cabi.PREC_FP32, "bf16": _cabi.PREC_BF16, "fp32tc": _cabi.PREC_FP32_TC}[prec]
        if prec == "fp32tc" and not self.fp32_tc:
            self.fp32_tc = True           # packs the extra blobs on this (first) use
        dev = asr.device
        f32 = lambda t: None if t is None else t.to(dev, torch.float32).contiguous()
        asr_, f0_, n_, s_ = f32(asr), f32(F0_curve), f32(N), f32(s)
        noise_, hs_, har_ = f32(noise), f32(har_source), f32(har)
        spf = self.samples_per_frame
        Lw = T * spf
        if noise_ is not None and tuple(noise_.shape) != (B, Lw, 9):
            raise RuntimeError(f"noise must be [B, {Lw}, 9]")
        if hs_ is not None and tuple(hs_.shape) != (B, Lw):
            raise RuntimeError(f"har_source must be [B, {Lw}]")
        if har_ is not None and (self.KIND != _cabi.KIND_ISTFTNET or tuple(har_.shape) != (B, 22, Lw // 5 + 1)):
            raise RuntimeError("har must be [B, 22, 60T+1] (istftnet only)")
        h = self._handle_for(dev)
        with torch.cuda.device(dev):
            need = lib.st2_decoder_workspace_bytes(h, B, T, prec_id)
            if need < 0:
                check(int(need), "st2_decoder_workspace_bytes")
            limit = int(self.workspace_limit_bytes)
            if limit and need > limit and B > 1 and not (return_excitation or return_taps):
                # micro-batches: rows are independent, results are identical to the one-shot call
                mb = max(1, min(B - 1, int(B * limit // need)))
                while mb > 1 and lib.st2_decoder_workspace_bytes(h, mb, T, prec_id) > limit:
                    mb -= 1
                wav = torch.empty(B, 1, Lw, dtype=torch.float32, device=dev)
                seed0, launches = self.seed, 0
                sl = lambda t, a, b: None if t is None else t[a:b]
                graphs, self.cuda_graphs = self.cuda_graphs, False  # a graph would pin one workspace per shape
                try:
                    for lo in range(0, B, mb):
                        hi = min(B, lo + mb)
                        self.seed = seed0
                        wav[lo:hi] = self._forward_native(
                            asr_[lo:hi], f0_[lo:hi], n_[lo:hi], s_[lo:hi], noise=sl(noise_, lo, hi),
                            har_source=sl(hs_, lo, hi), har=sl(har_, lo, hi), precision=prec,
                            item_offset=item_offset + lo, alignment=sl(alignment, lo, hi))
                        launches += self.last_launches
                finally:
                    self.cuda_graphs = graphs
                self.last_launches = launches
                return wav
            if (self.cuda_graphs and noise_ is None and hs_ is None and har_ is None and alignment is None
                    and not (return_excitation or return_taps) and not torch.cuda.is_current_stream_capturing()):
                return self._forward_graph(h, asr_, f0_, n_, s_, int(need), prec_id, int(item_offset))
            ws = torch.empty(int(need), dtype=torch.uint8, device=dev)
            wav = torch.empty(B, 1, Lw, dtype=torch.float32, device=dev)
            opts = _cabi.St2ForwardOpts()
            opts.precision = prec_id
            opts.item_offset = int(item_offset)
            if alignment is not None:
                opts.align_idx, opts.align_len = alignment.data_ptr(), int(align_len)
            opts.noise, opts.har_source, opts.har = (t.data_ptr() if t is not None else None
                                                     for t in (noise_, hs_, har_))
            opts.seed = int(self.seed)
            exc = {}
            if return_excitation:
                if har_ is None:
                    exc["har_source"] = torch.empty(B, Lw, dtype=torch.float32, device=dev)
                    opts.har_source_out = exc["har_source"].data_ptr()
                if self.KIND == _cabi.KIND_ISTFTNET:
                    exc["har"] = torch.empty(B, 22, Lw // 5 + 1, dtype=torch.float32, device=dev)
                    opts.har_out = exc["har"].data_ptr()
            taps = {}
            if return_taps:
                taps["decode3"] = torch.empty(B, 512, T, dtype=torch.float32, device=dev)
                opts.tap_decode = taps["decode3"].data_ptr()
                L, ch = T, self._cfg.upsample_initial_channel
                for i in range(self._cfg.n_ups):
                    L, ch = L * self._cfg.upsample_rates[i], ch // 2
                    if self.KIND == _cabi.KIND_ISTFTNET and i == self._cfg.n_ups - 1:
                        L += 1
                    taps[f"stage{i}"] = torch.empty(B, ch, L, dtype=torch.float32, device=dev)
                    opts.tap_stage[i] = taps[f"stage{i}"].data_ptr()
            stream = torch.cuda.current_stream(dev).cuda_stream
            check(lib.st2_decoder_forward(h, ptr(asr_), ptr(f0_), ptr(n_), ptr(s_), ptr(wav), ptr(ws),
                                          int(need), B, T, C.byref(opts), C.c_void_p(stream)),
                  "st2_decoder_forward")
            self.seed += 1 if noise_ is None and hs_ is None and har_ is None else 0
            self.last_launches = int(lib.st2_decoder_last_launches(h))
            # ws / inputs are only recycled by torch's caching allocator in stream order
        if return_taps:
            return wav, taps
        if return_excitation:
            return wav, exc
        return wav


def _graph_forward_impl(self, h, asr_, f0_, n_, s_, need, prec_id, item_offset):
    """CUDA-graph replay of one forward (in-kernel Philox noise, no taps).  The graph owns static copies
    of the inputs, the workspace, the output and a device-side seed; weights are baked in, so a reload
    (new handle version) invalidates the cache."""
    dev = asr_.device
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    B, _, T2 = asr_.shape
    T = 2 * T2
    Lw = T * self.samples_per_frame
    key = (idx, B, T, prec_id, item_offset, self._native_version.get(idx))
    g = self._graphs.get(key)
    if g is None:
        for k in [k for k in self._graphs if k[0] == idx and k[-1] != key[-1]]:
            del self._graphs[k]          # stale weights
        while len(self._graphs) >= max(1, int(self.max_graphs)):
            del self._graphs[next(iter(self._graphs))]
        g = {"asr": torch.empty_like(asr_), "f0": torch.empty_like(f0_), "n": torch.empty_like(n_),
             "s": torch.empty_like(s_), "ws": torch.empty(need, dtype=torch.uint8, device=dev),
             "wav": torch.empty(B, 1, Lw, dtype=torch.float32, device=dev),
             "seed": torch.zeros(1, dtype=torch.int64, device=dev)}
        opts = _cabi.St2ForwardOpts()
        opts.precision = prec_id
        opts.item_offset = item_offset
        opts.seed_dev = g["seed"].data_ptr()
        g["opts"] = opts

        def launch():
            stream = torch.cuda.current_stream(dev).cuda_stream
            check(lib.st2_decoder_forward(h, ptr(g["asr"]), ptr(g["f0"]), ptr(g["n"]), ptr(g["s"]), ptr(g["wav"]),
                                          ptr(g["ws"]), need, B, T, C.byref(opts), C.c_void_p(stream)),
                  "st2_decoder_forward")
        for src, dst in ((asr_, g["asr"]), (f0_, g["f0"]), (n_, g["n"]), (s_, g["s"])):
            dst.copy_(src)
        launch()                          # eager warm-up: first-use allocations happen outside the capture
        torch.cuda.current_stream(dev).synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            launch()
        g["graph"] = graph
        g["launches"] = int(lib.st2_decoder_last_launches(h))
        self._graphs[key] = g
    for src, dst in ((asr_, g["asr"]), (f0_, g["f0"]), (n_, g["n"]), (s_, g["s"])):
        dst.copy_(src, non_blocking=True)
    g["seed"].fill_(int(self.seed))
    g["graph"].replay()
    self.seed += 1
    self.last_launches = g["launches"]
    return g["wav"].clone()


_DecoderBase._forward_graph = _graph_forward_impl


class IstftnetDecoder(_DecoderBase):
    """``Modules.istftnet.Decoder`` (istftnet.py:466-527)."""
    KIND = _cabi.KIND_ISTFTNET

    def __init__(self, dim_in=512, F0_channel=512, style_dim=64, dim_out=80,
                 resblock_kernel_sizes=[3, 7, 11], upsample_rates=[10, 6], upsample_initial_channel=512,
                 resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]], upsample_kernel_sizes=[20, 12],
                 gen_istft_n_fft=20, gen_istft_hop_size=5):
        super().__init__()
        self._setup(_mk_config(self.KIND, dim_in, style_dim, dim_out, resblock_kernel_sizes, upsample_rates,
                               upsample_initial_channel, resblock_dilation_sizes, upsample_kernel_sizes,
                               gen_istft_n_fft, gen_istft_hop_size))


class HifiganDecoder(_DecoderBase):
    """``Modules.hifigan.Decoder`` (hifigan.py:415-474)."""
    KIND = _cabi.KIND_HIFIGAN

    def __init__(self, dim_in=512, F0_channel=512, style_dim=64, dim_out=80,
                 resblock_kernel_sizes=[3, 7, 11], upsample_rates=[10, 5, 3, 2], upsample_initial_channel=512,
                 resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]],
                 upsample_kernel_sizes=[20, 10, 6, 4]):
        super().__init__()
        self._setup(_mk_config(self.KIND, dim_in, style_dim, dim_out, resblock_kernel_sizes, upsample_rates,
                               upsample_initial_channel, resblock_dilation_sizes, upsample_kernel_sizes))

#!/usr/bin/env python
"""bench.py — audio-seconds decoded per second by the iSTFTNet waveform decoder on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|...]

One "step" = one Decoder.forward over one batch of synthetic utterances (BASELINE.json config 1:
LJSpeech iSTFTNet decoder, batch 32 x 800 mel frames = 320 s of 24 kHz audio, bf16 tensor-core path,
random-init weights).  Under torchrun every rank decodes its own batch of that size (pure data
parallelism, no collective on the path: SURVEY.md section 8e) -> "scaling": "weak"; `value` is the
whole job's audio-seconds divided by the max-over-ranks device time.

Prints ONE JSON line on rank 0 (contract in the task statement): value (inputs resident in HBM),
e2e (through the public module with pinned HOST buffers, H2D + D2H inside the timed region),
roofline (tcgen05 conv launches timed live with CUDA events on their stream), cpu_baseline (the
oracle port of the reference's CPU decoder on this box's host cores, bounded sample), clocks.

--impl reference times the reference's own CPU implementation of the path: the reference is pure
Python/PyTorch and /root/reference does not exist on the GPU box, so this arm runs the oracle port
(oracle/decoder_oracle.py: the same ATen CPU convolutions, same op order) with all host threads.

--impl reference-gpu times the LIBRARY baseline on the same B200:
the reference's op sequence (oracle port in library mode: F.conv1d / F.conv_transpose1d / F.instance_norm /
torch.stft / torch.istft on cuDNN, cuFFT and ATen) at the full workload, (i) eager fp32 with TF32 off
(train_second.py:183-186) and (ii) under torch.autocast('cuda', bfloat16).  The default line of the c2 / c4
workloads carries the same two figures in `library_baseline` (fewer steps).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (kind, B, T, description)   T = mel frames (hop 300 @ 24 kHz) -> 300*T samples
    "c1": ("istftnet", 1, 400, "BASELINE config 0: iSTFTNet B=1 x 400 frames (5 s)"),
    "c2": ("istftnet", 32, 800, "BASELINE config 1: LJSpeech iSTFTNet B=32 x 800 frames (10 s each)"),
    "c3": ("istftnet", 64, 800, "BASELINE config 2: multispeaker B=64 x 800 frames, one style row per item"),
    "c4": ("hifigan", 32, 800, "BASELINE config 3: HiFi-GAN decoder B=32 x 800 frames"),
    "c5": ("istftnet", 512, 2400, "BASELINE config 4 on one GPU: B=512 x 2400 frames, decoded in micro-batches"),
    "c5shard": ("istftnet", 64, 2400, "BASELINE config 4 per-GPU shard at 8 GPUs: B=64 x 2400 frames (30 s each)"),
}
DEFAULT_STEPS = {"c1": 500, "c2": 100, "c3": 50, "c4": 50, "c5": 2, "c5shard": 10, "f0n": 200, "sampler": 100, "chain": 40}
MFLOP_PER_FRAME = {"istftnet": 649.05, "hifigan": 873.90}  # SURVEY.md section 8d (2*MAC, convs + linears)


def _rows_of_rank(workload, B, world, rank):
    """(rows this rank decodes, global index of its first row, scaling).  Every workload is weak-scaled (each rank
    decodes its own batch of B rows) except c5: BASELINE.json config 4 is ONE batch of 512 x 2400 frames "sharded at
    1/2/4/8 B200", so its 512 rows are split over the ranks (strong scaling, styletts2_b200.shard.shard_range)."""
    if workload != "c5":
        return B, rank * B, "weak"
    from styletts2_b200.shard import shard_range
    lo, hi = shard_range(B, world, rank)
    return hi - lo, lo, "strong"


def _config(workload, world, graphs=True):
    """the `config` object of the JSON line; identical in the `ours` and `reference` arms (the reference arm's
    bounded sample is described in its cpu_baseline.sample)."""
    kind, B, T, desc = WORKLOADS[workload]
    act = B * (60 * T + 1) * 128 * 2 if kind == "istftnet" else B * 120 * T * 64 * 2
    per_rank = B if workload != "c5" else -(-B // world)
    big = act * (per_rank / B) > 126e6
    return {"workload": desc, "batch_per_gpu": per_rank, "mel_frames": T, "decoder": kind,
            "weights": "random-init, the reference constructor's distributions (synthetic.py init=\"reference\", seed 0)",
            "noise": "in-kernel Philox",
            "l2": (f"no explicit flush: one layer's activation is {act * (per_rank / B) / 1e9:.2f} GB per GPU (>> 126 MB L2) and "
                   "every step streams ~70 of them through HBM") if big else
                  (f"no explicit flush, and the working set is L2-sized (one layer's activation: {act * (per_rank / B) / 1e6:.1f} MB "
                   "against 126 MB of L2): a latency workload, replays run L2-warm as a real single-utterance loop would"),
            "cuda_graph": bool(graphs),
            "parallelism": (f"dp{world} (independent batches, no collective)" if workload != "c5" else
                            f"dp{world}: the 512 rows split into {per_rank} per GPU, no collective (strong scaling)")}
METRIC = "audio-sec/sec decoded (iSTFTNet, bf16)"


def _metric(kind):
    """BASELINE.json's metric for the iSTFTNet shapes; the HiFi-GAN shape (config 3) names its decoder"""
    return METRIC if kind == "istftnet" else "audio-sec/sec decoded (HiFi-GAN, bf16)"
UNIT = "audio-sec/sec"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(tf_sustained=d.get("bf16_tflops_sustained", 1357.0), tf_burst=d.get("bf16_tflops", 1628.5),
                    hbm=d.get("hbm_gbs", 6537.3), source="MEASURED_PEAKS.json")
    return dict(tf_sustained=1400.0, tf_burst=1590.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU during the timed region (NVML)."""

    def __init__(self, index: int, period=0.1):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = repr(e)

    def run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def _oracle_cpu_decoder(kind, B, T, steps, warmup, threads=None):
    """times the oracle port (= the reference's CPU algorithm) on the host cores; returns audio-s/s."""
    import torch
    from oracle import decoder_oracle as orc
    import synthetic as synth
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    cfg = synth.ISTFTNET_CFG if kind == "istftnet" else synth.HIFIGAN_CFG
    sd = synth.synth_state_dict(cfg, seed=0)
    inp = synth.synth_inputs(cfg, B, T, seed=1234)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        out = orc.decoder_forward(sd, cfg, inp["asr"], inp["F0_curve"], inp["N"], inp["s"], noise=inp["noise"])
        dt = time.perf_counter() - t0
        assert out.shape[-1] == 300 * T
        if i >= warmup:
            times.append(dt)
    audio_s = B * T * 300 / 24000.0
    return audio_s / statistics.median(times), statistics.median(times) * 1e3, torch.get_num_threads()


def _parity_paths(kind, cfg, T, dev, Bs=4):
    """Throughput of the two fp32 parity paths on a bounded sample of the workload (Bs items x T frames, device-resident,
    eager launches, CUDA events): "fp32" = every conv on the FFMA kernel (conv_simt), "fp32tc" = the dense convs on
    tcgen05 as three bf16 products with fp32 accumulation (ST2_PREC_FP32_TC).  Both meet the <= 1e-3 parity bar
    (tests/test_decoder_gpu.py); the headline runs the bf16 path."""
    import torch
    import styletts2_b200 as st2
    import synthetic as synth
    out = {"sample": f"{Bs} items x {T} frames of the workload, eager launches", "unit": UNIT}
    d = {k: v.to(dev) for k, v in synth.synth_inputs(cfg, Bs, T, seed=99, with_noise=False).items()}
    audio_s = Bs * T * 300 / 24000.0
    for prec, steps in (("fp32tc", 3), ("fp32", 1)):
        dec = (st2.IstftnetDecoder if kind == "istftnet" else st2.HifiganDecoder)(**{k: v for k, v in cfg.items() if k != "kind"})
        dec.load_state_dict(synth.synth_state_dict(cfg, seed=0), strict=True)
        dec = dec.to(dev).eval()
        dec.fp32_tc = prec == "fp32tc"
        try:
            dec(d["asr"], d["F0_curve"], d["N"], d["s"], precision=prec)   # warm-up (packs / uploads the weights)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                dec(d["asr"], d["F0_curve"], d["N"], d["s"], precision=prec)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            out[prec] = {"value": round(audio_s / (ms * 1e-3), 1), "ms_per_step": round(ms, 2), "launches": dec.last_launches}
        except RuntimeError as e:  # e.g. not enough memory next to the main arm's graph buffers
            out[prec] = {"unavailable": str(e)[:120]}
        del dec
        torch.cuda.empty_cache()
    return out


def _quality(kind, dev, dec, T, seed):
    """parity of the MEASURED build and module on ONE FULL-SIZE item of the measured workload (T mel frames, the
    reference constructor's init regime) against the oracle port of the reference, noise injected so both sides see
    the same excitation: the product chain in bf16 (phase scan -> har_source -> stft_hilo -> tcgen05 convs -> iSTFT):
    har_source error, torch.angle branch flips of `har`, waveform SNR (north star: >= 30 dB) and log-mel L1 with the
    reference's mel definition (meldataset.py:35-43); fp32 path max-abs (<= 1e-3).  The waveform comparisons feed the
    oracle with the har this run produced (iSTFTNet), which keeps the branch flips out of them."""
    import torch
    import synthetic as synth
    from oracle import decoder_oracle as orc
    cfg = synth.ISTFTNET_CFG if kind == "istftnet" else synth.HIFIGAN_CFG
    sd = synth.synth_state_dict(cfg, seed=0)
    inp = synth.synth_inputs(cfg, 1, T, seed=seed)
    graphs, dec.cuda_graphs = dec.cuda_graphs, False
    try:
        args = [inp[k].to(dev) for k in ("asr", "F0_curve", "N", "s")]
        w16, exc = dec(*args, noise=inp["noise"].to(dev), precision="bf16", return_excitation=True)
        hs = orc.source_module(inp["F0_curve"], inp["noise"], orc.Params(sd), 300)
        out = {"case": f"1 item x {T} frames of the workload (seed {seed}), noise injected, vs the oracle port of the reference",
               "har_source_max_abs": float((exc["har_source"].cpu() - hs).abs().max())}
        if kind == "istftnet":
            mag, ph = orc.stft_transform(hs, 20, 5)
            har = exc["har"].cpu()
            out["har_angle_branch_flips"] = int(((har[:, 11:] - ph).abs() > 3.0).sum())
            out["har_mag_max_abs"] = float((har[:, :11] - mag).abs().max())
            kw = {"har": har}
        else:
            kw = {"har_source": exc["har_source"].cpu()}
        ref = orc.decoder_forward(sd, cfg, inp["asr"], inp["F0_curve"], inp["N"], inp["s"], **kw)
        w32 = dec(*args, precision="fp32", **{k: v.to(dev) for k, v in kw.items()}).cpu()
        out.update({"fp32_max_abs": float((w32 - ref).abs().max()),
                    "bf16_snr_db": round(float(orc.snr_db(ref, w16.cpu())), 2),
                    "bf16_log_mel_l1": round(orc.log_mel_l1(ref, w16.cpu()), 5),
                    "ref_rms": round(float(ref.pow(2).mean().sqrt()), 4)})
    finally:
        dec.cuda_graphs = graphs
    return out


def _library_gpu(kind, B, T, steps, warmup, dev):
    """the library baseline on this GPU (SURVEY 8d / BASELINE.md section 4): the reference's op sequence on ATen
    (oracle port in library mode, oracle/decoder_oracle.py) at the full workload shape, the randn draw of SineGen
    inside the step like the reference; (i) eager fp32, TF32 off (train_second.py:183-186); (ii) bf16 autocast.
    Device-timed with CUDA events.  Falls back to micro-batches of 8 items when the eager graph does not fit."""
    import contextlib
    import torch
    import synthetic as synth
    from oracle import decoder_oracle as orc
    cfg = synth.ISTFTNET_CFG if kind == "istftnet" else synth.HIFIGAN_CFG
    sd = {k: v.to(dev) for k, v in synth.synth_state_dict(cfg, seed=0).items()}
    inp = {k: v.to(dev) for k, v in synth.synth_inputs(cfg, B, T, seed=1234, with_noise=False).items()}
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    audio_s = B * T * 300 / 24000.0
    out = {"what": "the reference's op sequence on cuDNN / cuFFT / ATen (oracle port, library mode) on this GPU, same "
                   f"workload (B={B} x {T} frames), randn noise drawn per step, eager launches",
           "unit": UNIT, "steps": steps, "warmup": warmup}
    mb = B
    for name, ctx in (("fp32_eager_tf32_off", contextlib.nullcontext),
                      ("bf16_autocast", lambda: torch.autocast("cuda", dtype=torch.bfloat16))):
        def step():
            for lo in range(0, B, mb):
                sl = slice(lo, min(B, lo + mb))
                noise = torch.randn(sl.stop - sl.start, T * 300, 9, device=dev)
                with ctx():
                    w = orc.decoder_forward(sd, cfg, inp["asr"][sl], inp["F0_curve"][sl], inp["N"][sl], inp["s"][sl],
                                            noise=noise, library=True)
            return w
        try:
            for _ in range(warmup):
                step()
        except torch.OutOfMemoryError:
            torch.cuda.empty_cache()
            mb = 8
            for _ in range(warmup):
                step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            w = step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        assert torch.isfinite(w.float()).all()
        out[name] = {"value": round(audio_s / (ms * 1e-3), 2), "ms_per_step": round(ms, 2), "micro_batch": mb,
                     "peak_mem_gb": round(torch.cuda.max_memory_allocated(dev) / 1e9, 1)}
    return out


def _hbm_kernels(B, T, dev, peaks):
    """the memory-bound kernels of the path (north star: "report achieved HBM GB/s"), timed alone at the
    workload's shape with CUDA events; L2 is flushed (256 MB memset) before every timed launch."""
    import torch
    from styletts2_b200._cabi import lib, check
    Lw, L2 = T * 300, T * 60 + 1
    g = torch.Generator().manual_seed(3)
    f0 = (torch.rand(B, T, generator=g) * 300 + 80).to(dev)
    lw = torch.randn(9, generator=g).to(dev)
    ph = torch.empty(B * 9 * T, device=dev)
    hs = torch.empty(B, Lw, device=dev)
    rows_pad = (L2 + 5) // 6 * 6 + 12
    har16 = torch.empty(B, rows_pad, 64, device=dev, dtype=torch.bfloat16)
    x = (torch.randn(B, L2, 32, generator=g) * 0.3).to(dev)
    wav = torch.empty(B, Lw, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    cases = [
        ("phase_scan + har_source (SineGen + SourceModuleHnNSF, Philox noise)", B * Lw * 4 + B * T * 4,
         lambda: check(lib.st2_source_har(f0.data_ptr(), None, 7, lw.data_ptr(), 0.0, B, T, 300, ph.data_ptr(),
                                          hs.data_ptr(), st))),
        ("stft_hilo_kernel (20-point DFT, |X| and angle as bf16 hi/lo rows: the STFT of the bf16 path)",
         B * Lw * 4 + B * rows_pad * 128,
         lambda: check(lib.st2_stft_har16(hs.data_ptr(), B, Lw, rows_pad, har16.data_ptr(), st))),
        ("istft_kernel (exp/sin heads, inverse DFT, overlap-add)", B * L2 * 88 + B * Lw * 4,
         lambda: check(lib.st2_istft_head(x.data_ptr(), B, L2, wav.data_ptr(), st))),
    ]
    out = []
    for name, nbytes, fn in cases:
        fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ms = []
        for _ in range(5):
            flush.zero_()
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        t = statistics.median(ms)
        gbs = nbytes / (t * 1e-3) / 1e9
        out.append({"kernel": name, "algorithmic_bytes": nbytes, "us": round(t * 1e3, 1), "achieved_GBs": round(gbs, 1),
                    "frac_of_hbm_peak": round(gbs / peaks["hbm"], 4)})
        if name.startswith("phase_scan"):
            out[-1]["note"] = ("instruction-bound, not HBM-bound: 9 Philox / Box-Muller normals, 9 range-reduced sines and a "
                               "tanh per output sample (~650 instructions per 4 bytes written)")
    return out


def run_f0n(args):
    """next-row workload: ProsodyPredictor.F0Ntrain at the C2 shape (B=32, 400 prosody
    frames -> 800 F0/N frames = 10 s of audio per item).  Same line format as the decoder bench."""
    import torch
    import torch.distributed as dist
    import styletts2_b200 as st2
    import synthetic as synth
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: styletts2_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, T = 32, 400
    cfg = synth.F0N_CFG
    p = st2.F0NPredictor(**cfg)
    sd = synth.synth_f0n_state_dict(cfg, seed=0)
    p.load_state_dict(sd, strict=True)
    p = p.to(dev).eval()
    p.precision = args.precision
    inp = synth.synth_f0n_inputs(cfg, B, T, seed=4321 + rank)
    host = {k: v.pin_memory() for k, v in inp.items()}
    d = {k: v.to(dev) for k, v in inp.items()}
    audio_s = B * 2 * T * 300 / 24000.0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(3, args.warmup)):
        f0, n = p.F0Ntrain(d["x"], d["s"])
    barrier()
    launches = p.last_launches
    sampler = ClockSampler(local)
    if sampler.ok:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        f0, n = p.F0Ntrain(d["x"], d["s"])
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    assert torch.isfinite(f0).all() and torch.isfinite(n).all()
    out_host = torch.empty(2, B, 2 * T).pin_memory()

    def e2e_step():
        a = {k: host[k].to(dev, non_blocking=True) for k in ("x", "s")}
        f0, n = p.F0Ntrain(a["x"], a["s"])
        out_host[0].copy_(f0, non_blocking=True)
        out_host[1].copy_(n, non_blocking=True)

    for _ in range(2):
        e2e_step()
    barrier()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import predictor_oracle as porc
        torch.set_num_threads(os.cpu_count() or 1)
        one = synth.synth_f0n_inputs(cfg, 1, T, seed=1)
        ts = []
        for i in range(3):
            t0 = time.perf_counter()
            porc.f0n_forward(sd, one["x"], one["s"], aten_lstm=True)
            ts.append(time.perf_counter() - t0)
        tm = statistics.median(ts[1:])
        cpu = {"value": round(2 * T * 300 / 24000.0 / tm, 3), "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
               "sample": f"1 item x {T} prosody frames, fp32, torch CPU (oracle port of models.py F0Ntrain, recurrence on ATen's nn.LSTM like the reference), {tm * 1e3:.0f} ms"}
    # algorithmic work: LSTM 2 dirs x (4H x (In + H)) + the six AdainResBlk1d convs + projections, 2*MAC per prosody frame
    H, In, dd = 256, 640, 512
    lstm = 2 * 2 * 4 * H * (In + H)
    blk = lambda din, dout, up: 2 * 3 * (din * dout * (2 if up else 1) + dout * dout * (2 if up else 1)) + (2 * din * dout * (2 if up else 1) if din != dout else 0)
    per_frame = lstm + 2 * (blk(dd, dd, False) + blk(dd, H, True) + blk(H, H, False) * 2 + 2 * H * 2)
    line = {
        "metric": "audio-sec/sec of F0/N curves predicted (ProsodyPredictor.F0Ntrain, bf16)", "value": round(world * audio_s / (ms_step * 1e-3), 2),
        "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": round(ms_step, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": "next row N1: F0Ntrain, B=32 x 400 prosody frames (800 F0 frames, 10 s each)", "batch_per_gpu": B,
                   "frames": T, "weights": "random-init (synthetic.py, seed 0)", "l2": "working set 0.3 GB per step (> 126 MB L2)",
                   "parallelism": f"dp{world} (independent batches, no collective)"},
        "e2e": {"value": round(world * audio_s / (ms_e2e * 1e-3), 2), "unit": UNIT, "ms_per_step": round(ms_e2e, 3),
                "h2d_bytes_per_step": sum(host[k].numel() * 4 for k in host), "d2h_bytes_per_step": out_host.numel() * 4},
        "gpu_launches": int(launches * args.steps),
        "roofline": {"bound": "tensor", "achieved": round(per_frame * B * T / (ms_step * 1e-3) / 1e12, 2), "peak": _peaks()["tf_sustained"],
                     "unit": "TFLOP/s", "frac": round(per_frame * B * T / (ms_step * 1e-3) / 1e12 / _peaks()["tf_sustained"], 4),
                     "traffic": None, "note": "whole forward (the LSTM recurrence is latency-bound: T dependent steps)"},
        "cpu_baseline": cpu, "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


SAMPLER_SHAPE = dict(B=64, n=128, steps=5)  # BASELINE config 2: "style s from ADPM2 diffusion sampler, 5 steps", batch 64


def sampler_flops(cfg, B, n, steps, guided=False):
    """algorithmic 2*MAC of one sampling run: per denoiser call and token the five linears of every block
    (to_q, to_kv, to_out, feed_forward.0/.2) + attention (QK^T and PV over n keys); per call and item the
    mapping MLP / to_time / to_out; once per item the AdaLayerNorm fc and to_features (modules.py:40-278)."""
    C, S, E = cfg["channels"], cfg["context_features"], cfg["context_embedding_features"]
    F, mid, L, m = C + E, cfg["num_heads"] * cfg["head_features"], cfg["num_layers"], cfg["multiplier"]
    per_tok = L * (2 * F * mid + 2 * F * 2 * mid + 2 * mid * F + 2 * 2 * F * F * m + 2 * 2 * n * mid)
    per_item_call = 2 * 2 * F * F + 2 * F * C
    runs = 2 if guided else 1
    calls = 2 * (steps - 1)
    once = (2 * S * L * 4 * F + 2 * S * F) if S else 0
    return B * (calls * (runs * (n * per_tok + 2 * F * C) + per_item_call) + once) + calls * 2 * (C + 1) * F


def run_sampler(args):
    """next-row workload: the ADPM2 style sampler at BASELINE config 2's shape (B=64 items, 128 text
    tokens each, 5 diffusion steps = 8 denoiser calls, embedding_scale 1 like Inference_LibriTTS.ipynb cell 20)."""
    import torch
    import torch.distributed as dist
    import styletts2_b200 as st2
    import synthetic as synth
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: styletts2_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, n, steps = SAMPLER_SHAPE["B"], SAMPLER_SHAPE["n"], SAMPLER_SHAPE["steps"]
    cfg = synth.SAMPLER_CFG
    m = st2.StyleSampler(**cfg)
    sd = synth.synth_sampler_state_dict(cfg, seed=0)
    m.load_state_dict(sd, strict=True)
    m = m.to(dev).eval()
    m.precision = args.precision
    inp = synth.synth_sampler_inputs(cfg, B, n, steps, seed=4321 + rank)
    del inp["step_noise"]  # throughput runs draw the step noise in-kernel (Philox)
    host = {k: v.pin_memory() for k, v in inp.items()}
    d = {k: v.to(dev) for k, v in inp.items()}
    audio_s = B * 10.0  # every item conditions one 800-frame (10 s) utterance of config 2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    call = lambda a, i: m(a["noise"], num_steps=steps, embedding=a["embedding"], features=a["features"], seed=i,
                          item_offset=rank * B)
    for i in range(max(3, args.warmup)):
        out = call(d, i)
    barrier()
    launches = m.last_launches
    sampler = ClockSampler(local)
    if sampler.ok:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        out = call(d, 100 + i)
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    assert torch.isfinite(out).all()
    out_host = torch.empty(B, 1, cfg["channels"]).pin_memory()

    def e2e_step(i):
        a = {k: host[k].to(dev, non_blocking=True) for k in host}
        out_host.copy_(call(a, 200 + i), non_blocking=True)

    for i in range(2):
        e2e_step(i)
    barrier()
    e0.record()
    for i in range(args.steps):
        e2e_step(i)
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        v, tm, cores = _oracle_cpu_sampler(1, n, steps, 2, 1)
        cpu = {"value": round(v, 3), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"1 item x {n} tokens x {steps} steps, fp32, torch CPU (oracle port of DiffusionSampler/ADPM2), {tm:.0f} ms"}
    fl = sampler_flops(cfg, B, n, steps)
    tf = fl / (ms_step * 1e-3) / 1e12
    line = {
        "metric": "audio-sec/sec of style vectors sampled (ADPM2 sampler over StyleTransformer1d, bf16)",
        "value": round(world * audio_s / (ms_step * 1e-3), 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": f"next row N2: style sampler, B={B} items x {n} text tokens, {steps} diffusion steps "
                               f"({2 * (steps - 1)} denoiser calls), embedding_scale 1; one item = the style of one 10 s utterance",
                   "batch_per_gpu": B, "tokens": n, "diffusion_steps": steps, "items_per_sec": round(world * B / (ms_step * 1e-3), 1),
                   "weights": "random-init (synthetic.py, seed 0)", "noise": "in-kernel Philox",
                   "l2": "activations of one denoiser call: 0.2 GB (> 126 MB L2); the 36 MB of bf16 weights stay L2-resident by design",
                   "parallelism": f"dp{world} (independent batches, no collective)"},
        "e2e": {"value": round(world * audio_s / (ms_e2e * 1e-3), 2), "unit": UNIT, "ms_per_step": round(ms_e2e, 3),
                "h2d_bytes_per_step": sum(host[k].numel() * 4 for k in host), "d2h_bytes_per_step": out_host.numel() * 4},
        "gpu_launches": int(launches * args.steps),
        "roofline": {"bound": "tensor", "achieved": round(tf, 2), "peak": _peaks()["tf_sustained"], "unit": "TFLOP/s",
                     "frac": round(tf / _peaks()["tf_sustained"], 4), "traffic": None,
                     "note": "whole sampling run (8 dependent denoiser calls of 3 transformer blocks)"},
        "cpu_baseline": cpu, "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_chain(args):
    """BASELINE config 2 taken literally: "multispeaker decoder (style s from ADPM2 diffusion sampler, 5 steps) batch
    64 x 800 frames".  One step = sampler(noise, embedding, features) -> s_pred; F0Ntrain(en, s_pred[:, 128:]) ->
    F0, N; Decoder(asr, F0, N, s_pred[:, :128]) -> waveform (Demo/Inference_LibriTTS.ipynb cell 16 from the
    sampler call on), all three on this package's kernels, B=64 items, 128 text tokens, 400 prosody frames."""
    import torch
    import torch.distributed as dist
    import styletts2_b200 as st2
    import synthetic as synth
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: styletts2_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, T, n, steps = 64, 800, SAMPLER_SHAPE["n"], SAMPLER_SHAPE["steps"]
    smp = st2.StyleSampler(**synth.SAMPLER_CFG)
    smp.load_state_dict(synth.synth_sampler_state_dict(synth.SAMPLER_CFG, seed=0), strict=True)
    f0n = st2.F0NPredictor(**synth.F0N_CFG)
    f0n.load_state_dict(synth.synth_f0n_state_dict(synth.F0N_CFG, seed=0), strict=True)
    dcfg = synth.ISTFTNET_CFG
    dec = st2.IstftnetDecoder(**{k: v for k, v in dcfg.items() if k != "kind"})
    dec.load_state_dict(synth.synth_state_dict(dcfg, seed=0), strict=True)
    smp, f0n, dec = smp.to(dev).eval(), f0n.to(dev).eval(), dec.to(dev).eval()
    for m in (smp, f0n, dec):
        m.precision = args.precision
    dec.cuda_graphs = not args.no_graphs
    si = synth.synth_sampler_inputs(synth.SAMPLER_CFG, B, n, steps, seed=4321 + rank)
    di = synth.synth_inputs(dcfg, B, T, seed=1234 + rank, with_noise=False)
    fi = synth.synth_f0n_inputs(synth.F0N_CFG, B, T // 2, seed=99 + rank)
    host = {"noise": si["noise"], "embedding": si["embedding"], "features": si["features"], "en": fi["x"], "asr": di["asr"]}
    host = {k: v.contiguous().pin_memory() for k, v in host.items()}
    d = {k: v.to(dev) for k, v in host.items()}
    audio_s = B * T * 300 / 24000.0

    def step(a, i):
        s_pred = smp(a["noise"], num_steps=steps, embedding=a["embedding"], features=a["features"], seed=i,
                     item_offset=rank * B).squeeze(1)
        F0, N = f0n.F0Ntrain(a["en"], s_pred[:, 128:].contiguous())
        return dec(a["asr"], F0, N, s_pred[:, :128].contiguous())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for i in range(max(3, args.warmup)):
        wav = step(d, i)
    barrier()
    launches = smp.last_launches + f0n.last_launches + dec.last_launches
    sampler = ClockSampler(local)
    if sampler.ok:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        wav = step(d, 100 + i)
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    assert wav.shape == (B, 1, 300 * T) and torch.isfinite(wav).all()
    wav_host = torch.empty(B, 1, 300 * T).pin_memory()

    def e2e_step(i):
        a = {k: host[k].to(dev, non_blocking=True) for k in host}
        wav_host.copy_(step(a, 200 + i), non_blocking=True)

    for i in range(3):
        e2e_step(i)
    barrier()
    e0.record()
    for i in range(args.steps):
        e2e_step(i)
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    if rank == 0:
        fl = MFLOP_PER_FRAME["istftnet"] * 1e6 * B * T + sampler_flops(synth.SAMPLER_CFG, B, n, steps)
        tf = fl / (ms_step * 1e-3) / 1e12
        line = {
            "metric": METRIC, "value": round(world * audio_s / (ms_step * 1e-3), 2), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": round(ms_step, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": {"workload": "BASELINE config 2, literal: ADPM2 style sampler (5 steps, 128 tokens) -> F0Ntrain -> iSTFTNet "
                                   "decoder, B=64 x 800 frames, chained on the device",
                       "batch_per_gpu": B, "mel_frames": T, "tokens": n, "diffusion_steps": steps,
                       "weights": "random-init (synthetic.py, seed 0)", "noise": "in-kernel Philox",
                       "l2": "every step streams > 6 GB through HBM (>> 126 MB L2)", "cuda_graph": bool(dec.cuda_graphs),
                       "parallelism": f"dp{world} (independent batches, no collective)"},
            "e2e": {"value": round(world * audio_s / (ms_e2e * 1e-3), 2), "unit": UNIT, "ms_per_step": round(ms_e2e, 3),
                    "h2d_bytes_per_step": sum(v.numel() * 4 for v in host.values()), "d2h_bytes_per_step": wav_host.numel() * 4,
                    "api": "StyleSampler.forward -> F0NPredictor.F0Ntrain -> IstftnetDecoder.forward, pinned host in/out"},
            "gpu_launches": int(launches * args.steps),
            "roofline": {"bound": "tensor", "achieved": round(tf, 2), "peak": _peaks()["tf_sustained"], "unit": "TFLOP/s",
                         "frac": round(tf / _peaks()["tf_sustained"], 4), "traffic": None,
                         "note": "whole chain (decoder + sampler algorithmic FLOPs; F0Ntrain not counted)"},
            "cpu_baseline": None, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def _oracle_cpu_sampler(B, n, steps, reps, warm):
    import torch
    import synthetic as synth
    from oracle import sampler_oracle as sorc
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = synth.SAMPLER_CFG
    sd = synth.synth_sampler_state_dict(cfg, seed=0)
    one = synth.synth_sampler_inputs(cfg, B, n, steps, seed=1)
    ts = []
    for i in range(warm + reps):
        t0 = time.perf_counter()
        with torch.no_grad():
            sorc.sample(sd, cfg, one["noise"], one["embedding"], one["features"], steps, 1.0, one["step_noise"])
        ts.append(time.perf_counter() - t0)
    tm = statistics.median(ts[warm:])
    return B * 10.0 / tm, tm * 1e3, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if args.workload == "sampler":
        n, steps = SAMPLER_SHAPE["n"], SAMPLER_SHAPE["steps"]
        v, tm, cores = _oracle_cpu_sampler(1, n, steps, args.steps, max(1, args.warmup))
        print(json.dumps({"impl": "reference", "metric": "audio-sec/sec of style vectors sampled (ADPM2 sampler over StyleTransformer1d, bf16)",
                          "value": round(v, 3), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": max(1, args.warmup),
                          "ms_per_step": round(tm, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic", "config": {"workload": "next row N2: style sampler", "sample": f"1 item x {n} tokens x {steps} steps per step on the CPU"},
                          "cpu_baseline": {"value": round(v, 3), "unit": UNIT, "cores": cores, "kind": "port",
                                           "sample": f"1 item x {n} tokens x {steps} steps, fp32, torch CPU (oracle port of Modules/diffusion)"},
                          "e2e": {"value": round(v, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)
        return 0
    if args.workload == "f0n":
        import torch
        import synthetic as synth
        from oracle import predictor_oracle as porc
        torch.set_num_threads(os.cpu_count() or 1)
        sd = synth.synth_f0n_state_dict(synth.F0N_CFG, seed=0)
        one = synth.synth_f0n_inputs(synth.F0N_CFG, 1, 400, seed=1)
        ts = []
        for i in range(max(1, args.warmup) + args.steps):
            t0 = time.perf_counter()
            porc.f0n_forward(sd, one["x"], one["s"], aten_lstm=True)   # nn.LSTM, what models.py:562-564 runs
            ts.append(time.perf_counter() - t0)
        tm = statistics.median(ts[max(1, args.warmup):])
        v = 10.0 / tm
        print(json.dumps({"impl": "reference", "metric": "audio-sec/sec of F0/N curves predicted (ProsodyPredictor.F0Ntrain, bf16)",
                          "value": round(v, 3), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": max(1, args.warmup),
                          "ms_per_step": round(tm * 1e3, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "f32", "data": "synthetic",
                          "config": {"workload": "next row N1: F0Ntrain", "sample": "B=1 x 400 prosody frames per step on the CPU"},
                          "cpu_baseline": {"value": round(v, 3), "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                                           "sample": "1 item x 400 prosody frames, fp32, torch CPU (oracle port of models.py F0Ntrain)"},
                          "e2e": {"value": round(v, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}), flush=True)
        return 0
    kind, B, T, desc = WORKLOADS[args.workload]
    sb, st = 1, T  # bounded sample: one utterance of the workload per step
    v, ms, threads = _oracle_cpu_decoder(kind, sb, st, args.steps, max(1, args.warmup))
    line = {
        "impl": "reference", "metric": _metric(kind), "value": round(v, 4), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(1, args.warmup), "ms_per_step": round(ms, 2),
        "higher_is_better": True, "scaling": "strong" if args.workload == "c5" else "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": _config(args.workload, max(1, args.gpus), graphs=not args.no_graphs),
        "cpu_baseline": {"value": round(v, 4), "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"each step = {sb} utterance x {st} frames of the workload (its rows are independent, "
                                   f"so audio-sec/sec does not depend on the batch), fp32, torch CPU on all host threads "
                                   f"(oracle port of Modules/{kind}.py Decoder.forward; /root/reference is absent on the GPU box)"},
        "e2e": {"value": round(v, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def run_reference_gpu(args):
    """--impl reference-gpu: the library baseline (cuDNN / cuFFT / ATen) on the B200, full workload, N=1."""
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "reference-gpu", "unavailable": "no CUDA device"}), flush=True)
        return 0
    kind, B, T, desc = WORKLOADS[args.workload]
    if args.workload == "c5":
        B = 64  # one 8-GPU shard; 512 x 2400 eager fp32 does not fit
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    sampler = ClockSampler(0)
    if sampler.ok:
        sampler.start()
    lib_ = _library_gpu(kind, B, T, args.steps, max(3, args.warmup), dev)
    clocks = sampler.stop()
    best = lib_["bf16_autocast"]
    line = {"impl": "reference-gpu", "metric": _metric(kind), "value": best["value"], "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": best["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16 (autocast); fp32 figure in library_baseline", "data": "synthetic",
            "config": _config(args.workload, 1, graphs=False), "library_baseline": lib_, "clocks": clocks,
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def run_ours(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    import styletts2_b200 as st2
    from styletts2_b200 import _cabi
    import synthetic as synth  # random-init weights / seeded inputs (no decoder arithmetic)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: styletts2_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    kind, B_job, T, desc = WORKLOADS[args.workload]
    B, row0, scaling = _rows_of_rank(args.workload, B_job, world, rank)
    cfg = synth.ISTFTNET_CFG if kind == "istftnet" else synth.HIFIGAN_CFG
    prec = args.precision

    dec = (st2.IstftnetDecoder if kind == "istftnet" else st2.HifiganDecoder)(
        **{k: v for k, v in cfg.items() if k != "kind"})
    dec.load_state_dict(synth.synth_state_dict(cfg, seed=0), strict=True)
    dec = dec.to(dev).eval()
    dec.precision = prec
    dec.fp32_tc = prec == "fp32tc"
    dec.cuda_graphs = not args.no_graphs  # CUDA-graph replay of the forward is allowed
    inp = synth.synth_inputs(cfg, B, T, seed=1234 + rank, with_noise=False)  # per-rank shard, Philox noise
    host = {k: v.pin_memory() for k, v in inp.items()}
    d = {k: v.to(dev) for k, v in inp.items()}
    audio_s = B * T * 300 / 24000.0          # this rank's audio per step
    job_audio_s = (B_job if scaling == "strong" else world * B) * T * 300 / 24000.0
    fwd = lambda a: dec(a["asr"], a["F0_curve"], a["N"], a["s"], item_offset=row0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident arm
    for _ in range(max(3, args.warmup)):
        out = fwd(d)
    barrier()
    launches_per_step = dec.last_launches
    sampler = ClockSampler(local)
    if sampler.ok:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        out = fwd(d)
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_rank = e0.elapsed_time(e1) / args.steps
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    ms_step = ms_total / args.steps
    value = job_audio_s / (ms_step * 1e-3)
    assert torch.isfinite(out).all()
    per_rank_ms = [round(ms_rank, 3)]
    if world > 1:  # off the timed path: the spread behind the max-over-ranks figure
        t = torch.zeros(world, device=dev, dtype=torch.float64)
        t[rank] = ms_rank
        dist.all_reduce(t)
        per_rank_ms = [round(float(x), 3) for x in t.tolist()]

    # ---- roofline leg: the same K steps launched eagerly with a CUDA-event pair around every tcgen05
    # conv launch on the launch stream (a graph replay cannot carry the per-launch events)
    graphs_on = dec.cuda_graphs
    dec.cuda_graphs = False
    roof_steps = max(1, min(args.steps, 20))
    for _ in range(2):
        fwd(d)
    barrier()
    _cabi.lib.st2_prof_enable(1)
    _cabi.lib.st2_prof_read(None, None, None, 1)
    e0.record()
    for _ in range(roof_steps):
        fwd(d)
    e1.record()
    barrier()
    ms_step_eager = e0.elapsed_time(e1) / roof_steps
    umma_ms, umma_n, umma_fl = C.c_double(), C.c_int64(), C.c_double()
    _cabi.lib.st2_prof_read(C.byref(umma_ms), C.byref(umma_n), C.byref(umma_fl), 1)
    _cabi.lib.st2_prof_enable(0)
    dec.cuda_graphs = graphs_on

    # ---- end-to-end arm: pinned host inputs -> H2D -> forward -> D2H of the waveform, every step
    wav_host = torch.empty(B, 1, T * 300, dtype=torch.float32).pin_memory()
    h2d = sum(host[k].numel() * 4 for k in ("asr", "F0_curve", "N", "s"))
    d2h = wav_host.numel() * 4

    # the package's host-to-host loop: uploads / downloads of neighbouring steps overlap the forward
    from styletts2_b200.pipeline import HostPipeline
    pipe = HostPipeline(dec, dev, depth=2, item_offset=row0)
    for _ in range(max(3, args.warmup) + 1):  # warm-up: the caching allocator needs a few rounds until the waveform
        pipe.submit(host, wav_host)           # blocks handed to the download stream are recycled without cudaMalloc
    pipe.wait()
    barrier()
    e0.record()
    for _ in range(args.steps):
        pipe.submit(host, wav_host)
    e1.record(pipe.d2h)  # after the last download
    pipe.wait()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    e2e_value = job_audio_s / (ms_e2e * 1e-3)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks = _peaks()
    algo_tflop_step = MFLOP_PER_FRAME[kind] * 1e6 * B * T / 1e12
    roof = None
    if umma_n.value > 0 and umma_ms.value > 0:
        ach = umma_fl.value / (umma_ms.value * 1e-3) / 1e12
        traffic, traffic_src = None, None  # DRAM bytes per launch of this kernel, from the committed ncu capture (profiles/)
        tp = os.path.join(ROOT, "profiles", "conv_umma_traffic.json")
        if args.workload == "c2" and os.path.exists(tp):
            tj = json.load(open(tp))
            traffic, traffic_src = round(tj["dram_bytes_per_launch"]), tj.get("source")
        roof = {"bound": "tensor", "achieved": round(ach, 2), "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                "frac": round(ach / peaks["tf_sustained"], 4), "traffic": traffic,
                "traffic_unit": f"bytes per launch (dram read + write, ncu: {traffic_src})",
                "kernel": "conv_umma_kernel (tcgen05 implicit-GEMM conv)",
                "launches_per_step": umma_n.value // roof_steps,
                "kernel_ms_per_step": round(umma_ms.value / roof_steps, 3),
                "kernel_share_of_step": round(umma_ms.value / roof_steps / ms_step_eager, 4),
                "measured_in": f"a second pass of {roof_steps} of the same steps launched eagerly "
                               f"({ms_step_eager:.3f} ms per step) with CUDA events around each launch",
                "algo_tflop_per_step_in_kernel": round(umma_fl.value / roof_steps / 1e12, 4),
                "peak_source": peaks["source"] + " bf16_tflops_sustained (kernel timed inside a long step); "
                               f"burst peak {peaks['tf_burst']}",
                "whole_step_tflops": round(algo_tflop_step / (ms_step * 1e-3), 2)}
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        v, ms_cpu, threads = _oracle_cpu_decoder(kind, 1, T, steps=2, warmup=1)
        cpu = {"value": round(v, 4), "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"1 utterance x {T} frames of the workload, fp32, torch CPU, median of 2 "
                         f"({ms_cpu:.0f} ms per forward)"}
        if T <= 800:  # SURVEY 8d also asks for the single-thread figure (one more forward, a few seconds)
            v1, ms1, _ = _oracle_cpu_decoder(kind, 1, T, steps=1, warmup=0, threads=1)
            cpu["value_1_thread"] = round(v1, 4)
            cpu["sample"] += f"; 1 thread: {ms1:.0f} ms"
    ws_bytes = int(_cabi.lib.st2_decoder_workspace_bytes(dec._native[local], B, T,
                                                         {"bf16": _cabi.PREC_BF16, "fp32": _cabi.PREC_FP32,
                                                          "fp32tc": _cabi.PREC_FP32_TC}[prec]))
    line = {
        "metric": _metric(kind), "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": round(ms_step, 3), "higher_is_better": True,
        "scaling": scaling, "vs_baseline": None, "dtype": prec, "data": "synthetic",
        "config": _config(args.workload, world, graphs=dec.cuda_graphs),
        "workspace_gb": round(ws_bytes / 1e9, 2), "per_rank_ms_per_step": per_rank_ms,
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "styletts2_b200.pipeline.HostPipeline.submit (pinned host in -> pinned host out, copies of "
                       "neighbouring steps overlap the forward)",
                "ms_per_step": round(ms_e2e, 3)},
        "gpu_launches": int(launches_per_step * args.steps),
        "roofline": roof, "cpu_baseline": cpu, "clocks": clocks,
        "hbm_kernels": _hbm_kernels(min(B, 64), T, dev, peaks) if (world == 1 and kind == "istftnet") else None,
        "quality": _quality(kind, dev, dec, T, 1234) if world == 1 and not args.no_quality else None,
        "parity_paths": (_parity_paths(kind, cfg, T, dev) if world == 1 and args.workload in ("c2", "c4")
                         and prec == "bf16" and not args.no_parity_paths else None),
        "library_baseline": (_library_gpu(kind, B, T, 3, 2, dev)
                             if world == 1 and args.workload in ("c2", "c4") and not args.no_library_baseline else None),
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="timed steps (default per workload: a >= 2 s timed region)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-gpu"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS) + ["f0n", "sampler", "chain"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32", "fp32tc"])
    ap.add_argument("--no-parity-paths", action="store_true", help="skip timing the fp32 parity paths (FFMA and 3 x bf16)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-library-baseline", action="store_true", help="skip the cuDNN/ATen-on-this-GPU block (c2, c4)")
    ap.add_argument("--no-quality", action="store_true", help="skip the full-size parity item against the oracle")
    ap.add_argument("--no-graphs", action="store_true", help="launch every kernel eagerly instead of replaying a CUDA graph")
    args = ap.parse_args()
    if args.steps is None:
        args.steps = 20 if args.impl != "ours" else DEFAULT_STEPS[args.workload]
    if args.impl == "reference":
        return run_reference(args)
    if args.impl == "reference-gpu":
        return run_reference_gpu(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: relaunch under torchrun (the driver launches torchrun itself)
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    if args.workload == "sampler":
        return run_sampler(args)
    if args.workload == "chain":
        return run_chain(args)
    return run_f0n(args) if args.workload == "f0n" else run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

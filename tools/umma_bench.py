"""micro-benchmark of the tcgen05 conv kernel on the stage shapes of config C2 (CUDA events, warm).
usage: python tools/umma_bench.py [--one k,res,stats] [--iters N]"""
import ctypes as C
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from styletts2_b200 import _cabi
from styletts2_b200._cabi import lib, check


def make(B, L, cin, n, k, dil, act, with_res, with_stats, accumulate=False, f32out=False, ident=False):
    dev = "cuda"
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, L, cin, generator=g).to(torch.bfloat16).to(dev)
    w = (torch.randn(k, cin, n, generator=g) / math.sqrt(cin * k)).contiguous()
    elems = lib.st2_umma_packed_elems(k, cin, n)
    packed = torch.empty(elems, dtype=torch.int16)
    check(lib.st2_umma_pack_weights(w.data_ptr(), k, cin, cin, n, packed.data_ptr()))
    wd = packed.to(dev)
    bias = torch.randn(n, generator=g).to(dev)
    ss = torch.stack([1 + 0.1 * torch.randn(B, cin, generator=g), 0.1 * torch.randn(B, cin, generator=g)], -1).contiguous().to(dev)
    alpha = torch.ones(cin, device=dev)
    y = torch.zeros(B, L, n, device=dev, dtype=torch.float32 if f32out else torch.bfloat16)
    res = torch.randn(B, L, n, generator=g).to(torch.bfloat16).to(dev) if with_res else None
    tr = lib.st2_conv_tile_rows(1)
    nt = (L + tr - 1) // tr
    stats = torch.zeros(B, nt, n, 2, device=dev) if with_stats else None
    p = _cabi.St2ConvParams()
    p.B = B
    p.x = x.data_ptr(); p.x_dtype = _cabi.DT_BF16; p.x_bstride = L * cin; p.ldx = cin; p.Lin = L; p.cin = cin
    p.ss = None if ident else ss.data_ptr(); p.ss_stride = cin; p.act = act; p.slope = 0.1; p.alpha = alpha.data_ptr()
    p.k, p.dil, p.pad, p.stride = k, dil, dil * (k - 1) // 2, 1
    p.w = wd.data_ptr(); p.bias = bias.data_ptr(); p.n_cols = n
    p.y = y.data_ptr(); p.y_dtype = _cabi.DT_F32 if f32out else _cabi.DT_BF16; p.y_bstride = L * n; p.ldy = n; p.Lout = L
    if res is not None:
        p.res = res.data_ptr(); p.res_bstride = L * n; p.ldres = n
    p.out_scale = 1.0
    p.accumulate = 1 if accumulate else 0
    if stats is not None:
        p.stats = stats.data_ptr(); p.n_tiles = nt
    p.algo_flops = 2.0 * B * L * cin * n * k
    keep = (x, wd, bias, ss, alpha, y, res, stats)
    return p, keep


def time_it(p, iters):
    for _ in range(2):
        check(lib.st2_conv_umma(C.byref(p), None))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        check(lib.st2_conv_umma(C.byref(p), None))
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def time_multi(ps, iters):
    arr = (_cabi.St2ConvParams * len(ps))(*ps)
    for _ in range(2):
        check(lib.st2_conv_umma_multi(arr, len(ps), None))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        check(lib.st2_conv_umma_multi(arr, len(ps), None))
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def multi_main(iters):
    """fused (one kernel, CTAs alternate between the convs tile by tile) against separate launches"""
    B, S = 32, _cabi.ACT_SNAKE
    for tag, L, C in (("s2", 48001, 128), ("s1", 8000, 256)):
        for res in (False, True):
            made = [make(B, L, C, C, k, 1, S, res, True) for k in (3, 7, 11)]
            ps = [m[0] for m in made]
            sep = [time_it(p, iters) for p in ps]
            for combo in ((0, 2), (0, 1), (1, 2), (0, 1, 2), (0, 2, 1)):
                us = time_multi([ps[i] for i in combo], iters)
                ssum = sum(sep[i] for i in combo)
                fl = sum(ps[i].algo_flops for i in combo)
                ks = "+".join(f"k{(3, 7, 11)[i]}" for i in combo)
                print(f"{tag} {'c2(res)' if res else 'c1     '} {ks:11s} fused {us:8.1f} us  separate {ssum:8.1f} us  "
                      f"({100 * (us / ssum - 1):+5.1f} %)  {fl / us / 1e6:7.1f} TFLOP/s", flush=True)
            del made, ps
            torch.cuda.empty_cache()


def l2_main(iters):
    """would an item-group schedule (keep a few items' activations in the 126 MB L2) speed the memory-bound
    convs up?  The same stage-2 conv on 2 / 3 items (x + res + y = 74 / 111 MB: L2-resident when launched back to back)
    and on 32 items (1.2 GB: streams through HBM), microseconds per 148-tile round."""
    S = _cabi.ACT_SNAKE
    for k, res in ((3, False), (3, True), (7, True), (11, True)):
        for B in (2, 3, 32):
            p, keep = make(B, 48001, 128, 128, k, 1, S, res, True)
            p.B = B
            us = time_it(p, iters * (8 if B < 32 else 1))
            tiles = B * 188
            rounds = (tiles + 147) // 148
            print(f"s2 k{k} {'c2(res)' if res else 'c1     '} B={B:2d}: {us:8.1f} us  {tiles:5d} tiles = {rounds:2d} rounds of 148, "
                  f"{us / rounds:6.2f} us per round, {us / B:7.2f} us per item", flush=True)
            del p, keep
            torch.cuda.empty_cache()


def main():
    iters = 5
    if "--iters" in sys.argv:
        iters = int(sys.argv[sys.argv.index("--iters") + 1])
    if "--multi" in sys.argv:
        return multi_main(iters)
    if "--l2" in sys.argv:
        return l2_main(iters)
    B = 32
    cases = []
    if "--one" in sys.argv:
        k, r, s = sys.argv[sys.argv.index("--one") + 1].split(",")
        cases = [("one", 48001, 128, 128, int(k), 1, _cabi.ACT_SNAKE, bool(int(r)), bool(int(s)), False, False)]
    else:
        for k in (3, 7, 11):
            cases.append((f"s2 k{k} d1 c1(stats)", 48001, 128, 128, k, 1, _cabi.ACT_SNAKE, False, True, False, False))
            cases.append((f"s2 k{k} d1 c2(res,stats)", 48001, 128, 128, k, 1, _cabi.ACT_SNAKE, True, True, False, False))
        cases.append(("s2 k11 d5 c1(stats)", 48001, 128, 128, 11, 5, _cabi.ACT_SNAKE, False, True, False, False))
        cases.append(("s2 k11 acc(res)", 48001, 128, 128, 11, 1, _cabi.ACT_SNAKE, True, False, True, False))
        cases.append(("s2 k3 none", 48001, 128, 128, 3, 1, _cabi.ACT_SNAKE, False, False, False, False))
        cases.append(("s2 k3 lrelu none", 48001, 128, 128, 3, 1, _cabi.ACT_LRELU, False, False, False, False))
        cases.append(("s1 k3 c1", 8000, 256, 256, 3, 1, _cabi.ACT_SNAKE, False, True, False, False))
        cases.append(("s1 k7 c2", 8000, 256, 256, 7, 1, _cabi.ACT_SNAKE, True, True, False, False))
        cases.append(("s1 k11 c2", 8000, 256, 256, 11, 1, _cabi.ACT_SNAKE, True, True, False, False))
        cases.append(("conv_post k7 n32 f32", 48001, 128, 32, 7, 1, _cabi.ACT_LRELU, False, False, False, True))
    if "--case" in sys.argv:  # L,cin,n,k,dil,act(0 none,1 lrelu,2 snake),res,stats
        L_, ci, n_, k, d, a, r, st_ = (int(v) for v in sys.argv[sys.argv.index("--case") + 1].split(","))
        cases = [(f"case L{L_} c{ci} n{n_} k{k} d{d} act{a} res{r} stats{st_}", L_, ci, n_, k, d, a if a else -1, bool(r), bool(st_), False, False)]
    if "--hifi" in sys.argv:
        cases = []
        for k in (2, 5, 7):  # pair-form tap counts of the 32-channel stage (k = 3, 7, 11 -> 3, 5, 7 taps of 64 x 64)
            cases.append((f"s4 pair k{k} c1(stats)", 120000, 64, 64, k, 1, _cabi.ACT_SNAKE, False, True, False, False))
            cases.append((f"s4 pair k{k} c2(res,stats)", 120000, 64, 64, k, 1, _cabi.ACT_SNAKE, True, True, False, False))
        for k in (3, 7, 11):
            cases.append((f"s3 c64 k{k} c1(stats)", 120000, 64, 64, k, 1, _cabi.ACT_SNAKE, False, True, False, False))
        cases.append(("s3 c64 k1 ident", 120000, 64, 64, 1, 1, _cabi.ACT_NONE, False, False, False, False))
        cases.append(("s3 c64 k7 ident", 120000, 64, 64, 7, 1, _cabi.ACT_NONE, False, False, False, False))
        cases.append(("s3 c64 k7 ident stats", 120000, 64, 64, 7, 1, _cabi.ACT_NONE, False, True, False, False))
        cases.append(("s3 c64 k7 snake nostats", 120000, 64, 64, 7, 1, _cabi.ACT_SNAKE, False, False, False, False))
    if "--ident" in sys.argv:
        cases = [(f"s2 k{k} ident", 48001, 128, 128, k, 1, _cabi.ACT_NONE, False, False, False, False) for k in (1, 3, 7, 11)]
        cases += [(f"s2 k{k} ident stats", 48001, 128, 128, k, 1, _cabi.ACT_NONE, False, True, False, False) for k in (3,)]
        cases += [(f"s2 k{k} ss-only", 48001, 128, 128, k, 1, -1, False, False, False, False) for k in (3,)]
    for name, L, cin, n, k, dil, act, r, s, acc, f32 in cases:
        p, keep = make(B, L, cin, n, k, dil, max(act, 0), r, s, acc, f32, ident=(act == _cabi.ACT_NONE))
        us = time_it(p, iters)
        tf = p.algo_flops / us / 1e6
        gb = (B * L * cin * 2 + B * L * n * (4 if f32 else 2) * (1 + (1 if r else 0) + (1 if acc else 0))) / 1e9
        print(f"{name:28s} {us:9.1f} us  {tf:8.1f} TFLOP/s  {gb / (us * 1e-6):8.1f} GB/s (min traffic {gb:.2f} GB)")
        del keep
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

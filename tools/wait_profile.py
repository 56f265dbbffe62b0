"""Per-role barrier-wait accounting of conv_umma_kernel (ST2_UMMA_DBG bit 64, st2_umma_wait_profile) on the stage
shapes of config C2.  Prints, per case, the launch time, the SM clock implied by the MMA thread's cycle count and the
share of its loop each role spent waiting on each barrier.   usage: python tools/wait_profile.py [--hifi]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from styletts2_b200 import _cabi
from styletts2_b200._cabi import lib, check
from umma_bench import make, time_it

B = 32
S = _cabi.ACT_SNAKE
cases = [  # name, L, cin, n, k, dil, act, res, stats, acc
    ("s2 k3 c1", 48001, 128, 128, 3, 1, S, False, True, False),
    ("s2 k3 c2 res", 48001, 128, 128, 3, 1, S, True, True, False),
    ("s2 k7 c1", 48001, 128, 128, 7, 1, S, False, True, False),
    ("s2 k7 c2 res", 48001, 128, 128, 7, 1, S, True, True, False),
    ("s2 k11 c1", 48001, 128, 128, 11, 1, S, False, True, False),
    ("s2 k11 d5 c1", 48001, 128, 128, 11, 5, S, False, True, False),
    ("s2 k11 c2 res", 48001, 128, 128, 11, 1, S, True, True, False),
    ("s2 k11 c2 res+acc", 48001, 128, 128, 11, 1, S, True, False, True),
    ("s1 k3 c1", 8000, 256, 256, 3, 1, S, False, True, False),
    ("s1 k7 c2 res", 8000, 256, 256, 7, 1, S, True, True, False),
    ("s1 k11 c2 res", 8000, 256, 256, 11, 1, S, True, True, False),
]
if "--hifi" in sys.argv:
    cases = [
        ("s3 c64 k3 c1", 120000, 64, 64, 3, 1, S, False, True, False),
        ("s3 c64 k7 c1", 120000, 64, 64, 7, 1, S, False, True, False),
        ("s3 c64 k11 c1", 120000, 64, 64, 11, 1, S, False, True, False),
        ("s3 c64 k11 c2 res", 120000, 64, 64, 11, 1, S, True, True, False),
        ("s3 c64 k11 res+acc", 120000, 64, 64, 11, 1, S, True, False, True),
        ("s4 pair k2 c1", 120000, 64, 64, 2, 1, S, False, True, False),
        ("s4 pair k5 c1", 120000, 64, 64, 5, 1, S, False, True, False),
    ]
if "--multi" in sys.argv:
    from umma_bench import time_multi
    for tag, L, C_ in (("s2", 48001, 128), ("s1", 8000, 256)):
        for combo in ((3, 11), (3, 7, 11)):
            made = [make(B, L, C_, C_, k, 1, S, False, True) for k in combo]
            ps = [m[0] for m in made]
            arr = (_cabi.St2ConvParams * len(ps))(*ps)
            os.environ.pop("ST2_UMMA_DBG", None); check(lib.st2_umma_reload_env())
            us = time_multi(ps, 5)
            os.environ["ST2_UMMA_DBG"] = "64"; check(lib.st2_umma_reload_env())
            out = (C.c_uint64 * 16)()
            check(lib.st2_conv_umma_multi(arr, len(ps), None)); check(lib.st2_umma_wait_profile(out, 1))
            us_p = time_multi(ps, 3)
            check(lib.st2_umma_wait_profile(out, 1))
            v = [int(x) for x in out]
            pct = lambda i, base: 100.0 * v[i] / max(v[base], 1)
            print(f"{tag} fused {'+'.join('k%d' % k for k in combo):10s} {us:7.1f} us ({us_p:7.1f} instr.) cyc/tile {v[0] / max(v[4], 1):7.0f} | MMA waits: t_empty {pct(1, 0):4.1f}% "
                  f"a_full {pct(2, 0):4.1f}% w_full {pct(3, 0):4.1f}% | EPI: t_full {pct(6, 5):4.1f}% stg_free {pct(7, 5):4.1f}% stats {pct(8, 5):4.1f}% "
                  f"| PROD: a_empty {pct(10, 9):4.1f}% raw_full {pct(11, 9):4.1f}% math {pct(12, 9):4.1f}%", flush=True)
            os.environ.pop("ST2_UMMA_DBG", None); check(lib.st2_umma_reload_env())
            del made, ps, arr
            torch.cuda.empty_cache()
    sys.exit(0)
for name, L, cin, n, k, dil, act, r, s, acc in cases:
    p, keep = make(B, L, cin, n, k, dil, act, r, s, acc)
    os.environ.pop("ST2_UMMA_DBG", None)
    check(lib.st2_umma_reload_env())
    us = time_it(p, 5)
    os.environ["ST2_UMMA_DBG"] = "64"
    check(lib.st2_umma_reload_env())
    out = (C.c_uint64 * 16)()
    check(lib.st2_conv_umma(C.byref(p), None))
    check(lib.st2_umma_wait_profile(out, 1))
    us_p = time_it(p, 3)
    check(lib.st2_umma_wait_profile(out, 1))
    v = [int(x) for x in out]
    launches, ctas = 5, 148  # time_it: 2 warm + 3 timed
    loop = v[0] / launches / ctas
    mhz = loop / us_p
    tf = p.algo_flops / us / 1e6
    pct = lambda i, base: 100.0 * v[i] / max(v[base], 1)
    print(f"{name:20s} {us:7.1f} us ({us_p:7.1f} instrumented) {tf:7.1f} TF  ~{mhz:5.0f} MHz  tiles/CTA {v[4] / launches / ctas:5.1f} "
          f"cyc/tile {v[0] / max(v[4], 1):7.0f} | MMA waits: t_empty {pct(1, 0):4.1f}% a_full {pct(2, 0):4.1f}% w_full {pct(3, 0):4.1f}% "
          f"| EPI: t_full {pct(6, 5):4.1f}% stg_free {pct(7, 5):4.1f}% stats {pct(8, 5):4.1f}% "
          f"| PROD: a_empty {pct(10, 9):4.1f}% raw_full {pct(11, 9):4.1f}% math {pct(12, 9):4.1f}%", flush=True)
    os.environ.pop("ST2_UMMA_DBG", None)
    check(lib.st2_umma_reload_env())
    del keep
    torch.cuda.empty_cache()

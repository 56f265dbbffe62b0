"""CTA-pair mode of conv_umma (ST2_UMMA_PAIR=1) against the single-CTA kernel on the same inputs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from styletts2_b200 import _cabi
from util import run_conv


def case(B, L, cin, n, k, dil, act, res, stats, acc=False, f32=False, bias_on=True):
    g = torch.Generator().manual_seed(B * 1000 + L + k)
    x = torch.randn(B, L, cin, generator=g).to(torch.bfloat16).cuda()
    w = torch.randn(k, cin, n, generator=g) * (1.0 / (cin * k) ** 0.5)
    bias = torch.randn(n, generator=g) if bias_on else None
    ss = torch.stack([1 + 0.1 * torch.randn(B, cin, generator=g), 0.1 * torch.randn(B, cin, generator=g)], -1) if act != _cabi.ACT_NONE else None
    alpha = (0.5 + torch.rand(cin, generator=g)) if act == _cabi.ACT_SNAKE else None
    r = torch.randn(B, L, n, generator=g).to(torch.bfloat16) if res else None
    yo = torch.randn(B, L, n, generator=g).to(torch.bfloat16) if acc else None
    out = {}
    for mode in ("0", "1"):
        os.environ["ST2_UMMA_PAIR"] = mode
        _cabi.check(_cabi.lib.st2_umma_reload_env())
        try:
            y, st = run_conv("umma", x, w, bias, k=k, dil=dil, pad=dil * (k - 1) // 2, ss=ss, act=act, slope=0.1, alpha=alpha,
                             res=r, out_scale=0.5 if acc else 1.0, y_old=yo, want_stats=stats,
                             out_dtype=torch.float32 if f32 else None)
            torch.cuda.synchronize()
            out[mode] = (y.float().cpu(), None if st is None else st.float().cpu())
        except Exception as e:  # noqa: BLE001
            print("   mode", mode, "FAILED:", str(e)[:300])
            return False
    d = (out["0"][0] - out["1"][0]).abs().max().item()
    ds = 0.0 if out["0"][1] is None else (out["0"][1] - out["1"][1]).abs().max().item()
    ok = d == 0.0 and ds <= 1e-3 * max(1.0, out["0"][1].abs().max().item() if out["0"][1] is not None else 1.0)
    print(f"B{B} L{L} cin{cin} n{n} k{k} d{dil} act{act} res{int(res)} st{int(stats)} acc{int(acc)}: max|dy| {d:.3g} max|dstats| {ds:.3g} {'OK' if ok else 'MISMATCH'}")
    return ok


if __name__ == "__main__":
    S, L, N = _cabi.ACT_SNAKE, _cabi.ACT_LRELU, _cabi.ACT_NONE
    good = True
    for c in [(1, 256, 64, 128, 1, 1, N, False, False), (1, 512, 64, 128, 1, 1, N, False, False),
              (2, 700, 128, 128, 3, 1, S, False, True), (3, 1000, 128, 128, 7, 3, S, True, True),
              (2, 2000, 256, 256, 11, 5, S, True, True), (5, 777, 128, 128, 3, 1, L, False, True),
              (2, 1500, 128, 128, 7, 1, S, True, False, True), (40, 3000, 128, 128, 11, 1, S, True, True),
              (2, 700, 128, 32, 7, 1, L, False, False, False, True)]:
        good = case(*c) and good
    print("ALL OK" if good else "FAILURES")

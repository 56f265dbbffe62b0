"""Diagnostic for the tcgen05 conv: structured inputs that reveal descriptor / layout mistakes.
Prints, for a k=1 GEMM and a k=3 shifted conv, where the result deviates from the expectation."""
import sys, os, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from styletts2_b200 import _cabi
from tests.util import run_conv, conv_reference


def report(name, y, ref):
    y, ref = y.cpu().double(), ref.double()
    d = (y - ref).abs()
    print(f"[{name}] shape {tuple(y.shape)} finite={torch.isfinite(y).all().item()} max err {d.max().item():.4g} "
          f"ref absmax {ref.abs().max().item():.4g}")
    if d.max().item() > 0.05:
        bad = (d > 0.05)
        rows = bad.any(dim=2)[0]
        cols = bad.any(dim=1)[0]
        print("   bad rows (first 40):", torch.nonzero(rows).flatten()[:40].tolist())
        print("   bad cols (first 40):", torch.nonzero(cols).flatten()[:40].tolist())
        print("   y[0,:4,:8] =", y[0, :4, :8].tolist())
        print("   ref[0,:4,:8] =", ref[0, :4, :8].tolist())
        # does y equal ref under a row permutation within groups of 8 / a column shuffle?
        for r in range(0, 16):
            best = (y[0, r][None, :] - ref[0, :64]).abs().sum(1).argmin().item()
            print(f"   y row {r} is closest to ref row {best}")


def main():
    torch.manual_seed(0)
    # 1) k=1: y[t, n] = sum_c x[t, c] w[c, n]; x one-hot in c = t % 64, w[c, n] = c + n/256  -> y[t,n] = (t%64) + n/256
    Lin, cin, n = 256, 64, 128
    x = torch.zeros(1, Lin, cin)
    for t in range(Lin):
        x[0, t, t % 64] = 1.0
    w = torch.zeros(1, cin, n)
    for c in range(cin):
        for j in range(n):
            w[0, c, j] = c + j / 256.0
    ref = conv_reference(x, None, 0, 0.0, None, w.to(torch.bfloat16).float(), None, 1, 1, 0, 1, Lin, None, 0, 1.0, None)
    y, _ = run_conv("umma", x.to(torch.bfloat16).cuda(), w, None, k=1)
    report("k1 one-hot", y.float(), ref)
    # 2) k=3 dil=1: x[t, c] = delta(c == 0) * (t % 32);  w[j][0][n] = 10^j * delta... y[t, n=j] = x[t + j - 1]
    cin, n = 64, 128
    x = torch.zeros(1, Lin, cin)
    x[0, :, 0] = torch.arange(Lin) % 32
    w = torch.zeros(3, cin, n)
    for j in range(3):
        w[j, 0, j] = 1.0
    ref = conv_reference(x, None, 0, 0.0, None, w, None, 3, 1, 1, 1, Lin, None, 0, 1.0, None)
    y, _ = run_conv("umma", x.to(torch.bfloat16).cuda(), w, None, k=3, pad=1)
    report("k3 shift", y.float(), ref)
    # 3) random, two chunks, multi tile
    x = torch.randn(2, 700, 128)
    w = torch.randn(7, 128, 128) / 30
    ref = conv_reference(x.to(torch.bfloat16).float(), None, 0, 0.0, None, w.to(torch.bfloat16).float(), None, 7, 3, 9, 1, 700, None, 0, 1.0, None)
    y, _ = run_conv("umma", x.to(torch.bfloat16).cuda(), w, None, k=7, dil=3, pad=9)
    report("k7 d3 random", y.float(), ref)


if __name__ == "__main__":
    main()

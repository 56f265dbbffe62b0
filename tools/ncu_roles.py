"""split the SASS of conv_umma_kernel into warp roles by marker instructions and report, per role,
samples, executed warp-instructions and the share of samples spent spinning on mbarriers.
usage: python tools/ncu_roles.py src.csv"""
import csv, sys, re
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
S = lambda r: int(r[ix["# Samples"]] or 0)
E = lambda r: int(r[ix["Instructions Executed"]] or 0)
src = lambda r: r[ix["Source"]]
marks = []
for i, r in enumerate(body):
    s = src(r)
    for pat, role in (("UTCHMMA", "mma"), ("UBLKCP", "wprod"), ("LDGSTS", "prod"), ("LDTM", "epi"), ("UTMASTG", "epi")):
        if pat in s:
            marks.append((i, role))
# role of instruction i = role of nearest marker
import bisect
mi = [m[0] for m in marks]
def role_of(i):
    j = bisect.bisect_left(mi, i)
    c = []
    if j < len(mi): c.append((mi[j] - i, marks[j][1]))
    if j > 0: c.append((i - mi[j - 1], marks[j - 1][1]))
    return min(c)[1] if c else "?"
agg = {}
for i, r in enumerate(body):
    ro = role_of(i)
    a = agg.setdefault(ro, [0, 0, 0, 0])
    a[0] += S(r); a[1] += E(r)
    s = src(r)
    # spin: the TRYWAIT itself and the 3 instructions after it
    spin = any("TRYWAIT" in src(body[k]) for k in range(max(0, i - 3), i + 1))
    if spin: a[2] += S(r); a[3] += E(r)
tot = sum(a[0] for a in agg.values())
for ro, a in agg.items():
    print(f"{ro:6s} samples {a[0]:6d} ({100*a[0]/tot:4.1f}%)  warp-instr {a[1]:10d}  spin samples {a[2]:6d} ({100*a[2]/max(1,a[0]):4.1f}% of role)  spin instr {a[3]:9d}")
# per-barrier spin counts
print("barrier waits (executions of TRYWAIT by operand):")
for i, r in enumerate(body):
    if "TRYWAIT" in src(r):
        ss = sum(S(body[k]) for k in range(i, min(len(body), i + 4)))
        print(f"  {role_of(i):6s} {r[ix['Address']][-5:]} ex={E(r):8d} samples~{ss:5d}  {src(r).strip()[:80]}")

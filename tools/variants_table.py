"""table of gpurun_out/variants.log (gpurun_ab/run_variants.sh): one column per library variant"""
import re, collections, sys
rows = collections.OrderedDict(); cur = None; vs = []
for l in open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/variants.log"):
    if l.startswith("==="):
        cur = l.split()[1]; vs.append(cur); continue
    m = re.match(r"(.{28})\s+([\d.]+) us", l)
    if m:
        rows.setdefault(m.group(1).strip(), {})[cur] = m.group(2)
    elif l.split() and l.split()[0] == cur:
        rows.setdefault(" ".join(l.split()[1:2]), {})[cur] = "/".join(l.split()[2:5])
print(f"{'case':28s}" + "".join(f"{v:>10s}" for v in vs))
for k, d in rows.items():
    print(f"{k:28s}" + "".join(f"{d.get(v, ''):>10s}" if len(d.get(v, "")) < 10 else "  " + d[v] for v in vs))

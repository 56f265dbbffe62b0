"""top stall sites from an `ncu --page source --csv --print-source sass` dump.
usage: python tools/ncu_src_top.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
tot = sum(int(r[ix["# Samples"]] or 0) for r in body)
print("total samples", tot, "instructions", len(body))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(int(r[ix[s]] or 0) for r in body) for s in stalls}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
top = sorted(body, key=lambda r: -int(r[ix["# Samples"]] or 0))[:n]
for r in top:
    s = int(r[ix["# Samples"]])
    why = sorted(((int(r[ix[k]] or 0), k[6:]) for k in stalls), reverse=True)[:2]
    print(f"{r[ix['Address']][-5:]} {s:6d} {100*s/tot:5.1f}%  ex={r[ix['Instructions Executed']]:>9s}  {r[ix['Source']][:70]:70s} {why}")

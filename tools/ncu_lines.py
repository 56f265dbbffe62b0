"""per-CUDA-source-line aggregation of an `ncu -i rep --page source --csv --print-source cuda,sass` dump: samples,
executed warp instructions and the two dominant stall reasons per line, top N lines, plus totals per line range
(= warp role).  usage: python tools/ncu_lines.py srccs.csv [N]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
heads = [i for i, r in enumerate(rows) if len(r) > 3 and r[0] == "Line No" and r[2] == "Address"]
hdr = rows[heads[0]]
ix = {h: i for i, h in enumerate(hdr)}
ix["SASS"] = 3
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
per = collections.OrderedDict()
cur = None
for r in rows[heads[0] + 1:(heads[-1] if len(heads) > 1 else len(rows))]:
    if len(r) != len(hdr):
        continue
    if r[0] == "Line No":
        continue
    if r[0]:
        cur = int(r[0])
        per.setdefault(cur, {"src": r[1], "s": 0, "ex": 0, "st": collections.Counter(), "n": 0})
        continue
    if cur is None or r[2] in ("", "..."):
        continue
    try:
        s = int(r[ix["# Samples"]] or 0); ex = int(r[ix["Instructions Executed"]] or 0)
    except ValueError:
        continue
    d = per[cur]
    d["s"] += s; d["ex"] += ex; d["n"] += 1
    for k in stalls:
        try:
            d["st"][k[6:]] += int(r[ix[k]] or 0)
        except ValueError:
            pass
tot = sum(d["s"] for d in per.values()); totex = sum(d["ex"] for d in per.values())
print(f"total samples {tot}, executed warp instructions {totex}")
agg = collections.Counter()
for d in per.values():
    agg.update(d["st"])
print("stall reasons:", {k: v for k, v in agg.most_common() if v})
for ln, d in sorted(per.items(), key=lambda kv: -kv[1]["s"])[:n]:
    why = ", ".join(f"{k} {v}" for k, v in d["st"].most_common(3) if v)
    print(f"L{ln:5d} {d['s']:6d} {100 * d['s'] / max(tot, 1):5.1f}%  ex={d['ex']:9d} ({100 * d['ex'] / max(totex, 1):4.1f}%)  {d['src'].strip()[:78]:78s} | {why}")
# executed warp-instructions and samples per line range of conv_umma.cu (inlined helpers are attributed to their own lines)
if "--ranges" in sys.argv:
    R = [("ptx wrappers / barriers 129-260", 129, 260), ("mma asm 261-352", 261, 352), ("tmem ld 371-395", 371, 395),
         ("pack_bf16x2 412-415", 412, 415), ("f2 ops 417-436 (stats)", 417, 436), ("bf16lo/hi 437-438 (producer)", 437, 438),
         ("um_clk 449", 449, 449), ("tile bookkeeping 451-541", 451, 541), ("um_act 543-555", 543, 555),
         ("um_produce 557-849", 557, 849), ("um_stage32 858-896", 858, 896), ("um_epilogue 898-962", 898, 962),
         ("epi: staging 963-1003", 963, 1003), ("epi: stats 1004-1029", 1004, 1029), ("epi: sred 1030-1061", 1030, 1061),
         ("kernel prologue 1063-1160", 1063, 1160), ("mma warp 1161-1276", 1161, 1276), ("w producer 1277-1327", 1277, 1327),
         ("loader 1328-1375", 1328, 1375), ("tail 1376-1402", 1376, 1402)]
    print()
    for name, lo, hi in R:
        s = sum(d["s"] for ln, d in per.items() if lo <= ln <= hi); ex = sum(d["ex"] for ln, d in per.items() if lo <= ln <= hi)
        print(f"  {name:40s} samples {s:6d} ({100 * s / tot:4.1f}%)  warp-instr {ex:10d} ({100 * ex / totex:4.1f}%)")

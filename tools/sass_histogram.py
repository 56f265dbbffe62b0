"""SASS opcode histogram of the shipped library (cuobjdump -sass): per kernel, the Blackwell-specific opcodes that show
the tcgen05 / TMA / TMEM path is what was compiled.  usage: python tools/sass_histogram.py > profiles/<tag>_sass_histogram.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "styletts2_b200", "_lib", "libst2dec.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
WATCH = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "UTCATOM", "SYNCS", "HMMA", "FFMA2", "FADD2",
         "MUFU", "LDGSTS", "ELECT", "UCGABAR", "ACQBULK", "UTMACMDFLUSH"]
per = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        per[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        per[cur]["_total"] += 1
        for w in WATCH:
            if op == w or op.startswith(w + "."):
                per[cur][w] += 1
                if w in ("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "HMMA") and op != w:
                    per[cur]["  " + op] += 1
print(f"cuobjdump -sass {os.path.relpath(so, ROOT)} ({os.path.getsize(so)} bytes), sm_100a; opcode counts per kernel")
tot = collections.Counter()
for k, c in per.items():
    tot.update(c)
print("\nwhole library:")
for w, n in sorted(tot.items(), key=lambda kv: (-kv[1], kv[0])):
    if w != "_total":
        print(f"  {w:28s} {n}")
print(f"  instructions                 {tot['_total']}")
print("\nper kernel (kernels with at least one watched opcode):")
for k, c in per.items():
    hit = {w: n for w, n in c.items() if w != "_total"}
    if hit:
        print(f"  {k}  [{c['_total']} instr]")
        print("     " + ", ".join(f"{w} {n}" for w, n in sorted(hit.items())))

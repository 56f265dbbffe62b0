"""summarise an .ncu-rep (no GPU needed): key metrics, samples per warp role, top stall sites.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep"""
import csv, subprocess, sys, collections, io, re
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h, u, v = r[0], r[1], r[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.max", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum", "sm__warps_active.avg.pct_of_peak_sustained_active"]
for a, b, c in zip(h, u, v):
    if a in want:
        print(f"{a:85s} {c} {b}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hh = rows[1]; ix = {n: i for i, n in enumerate(hh)}; data = rows[2:]
def f(r, n):
    try: return float(r[ix[n]])
    except Exception: return 0.0
tots = sum(f(r, "# Samples") for r in data); toti = sum(f(r, "Instructions Executed") for r in data)
print("samples", tots, "inst", toti)
top = sorted(data, key=lambda r: -f(r, "# Samples"))[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]
for r in top:
    st = {k: int(f(r, k)) for k in ("stall_long_sb", "stall_barrier", "stall_short_sb", "stall_wait", "stall_mio", "stall_math", "stall_lg", "stall_branch_resolving", "stall_not_selected") if f(r, k) > 0}
    print(r[ix["Address"]][-5:], f'{f(r, "# Samples") / tots * 100:5.2f}%', int(f(r, "Instructions Executed")), st, r[ix["Source"]].strip()[:70])

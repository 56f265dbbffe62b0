"""turn the raw ncu CSVs of tools/gpu_profile.sh (gpurun_out/) into the committed summaries under profiles/.
usage: python tools/profiles_refresh.py <tag>   e.g. r1_v4"""
import collections, csv, json, os, subprocess, sys
tag = sys.argv[1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")

def rows_of(path):
    lines = [l for l in open(path) if l.startswith('"')]
    return list(csv.DictReader(lines))

# (1) launch list of one bench run -> per-kernel shares
agg = collections.OrderedDict()
for r in rows_of(os.path.join(G, "launches.csv")):
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"].split("(")[0]
    v = float(r["Metric Value"].replace(",", ""))
    v = v / 1e3 if r["Metric Unit"] in ("ns", "nsecond") else v  # -> us
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
with open(os.path.join(P, f"{tag}_launch_shares_c2.csv"), "w") as f:
    f.write("kernel,launches,total_ms,share_pct\n")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k},{a[0]},{a[1] / 1e3:.3f},{100 * a[1] / tot:.2f}\n")

# (2) per-launch metrics of every tcgen05 conv of one step
per = collections.OrderedDict()
for r in rows_of(os.path.join(G, "umma_step.csv")):
    per.setdefault(int(r["ID"]), {})[r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])
dram = 0.0
with open(os.path.join(P, f"{tag}_ncu_conv_umma_step_c2.csv"), "w") as f:
    f.write("id,time_us,dram_read_MB,dram_write_MB,tensor_active_pct,warp_inst_M,cycles\n")
    for i, m in per.items():
        t, u = m["gpu__time_duration.sum"]; t = t / 1e3 if u in ("ns", "nsecond") else t
        rd, wr = m["dram__bytes_read.sum"], m["dram__bytes_write.sum"]
        sc = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rdb, wrb = rd[0] * sc[rd[1]], wr[0] * sc[wr[1]]
        dram += rdb + wrb
        f.write(f"{i},{t:.1f},{rdb / 1e6:.1f},{wrb / 1e6:.1f},"
                f"{m['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'][0]:.1f},"
                f"{m['smsp__inst_executed.sum'][0] / 1e6:.1f},{int(m['sm__cycles_elapsed.max'][0])}\n")
json.dump({"kernel": "conv_umma_kernel", "launches_per_step": len(per), "dram_bytes_per_step": dram,
           "dram_bytes_per_launch": dram / max(1, len(per)),
           "source": f"profiles/{tag}_ncu_conv_umma_step_c2.csv (ncu, config C2, one step)"},
          open(os.path.join(P, "conv_umma_traffic.json"), "w"), indent=1)

# (3) the full capture: details page + top stall sites
rep = os.path.join(G, "prof_k11.ncu-rep")
if os.path.exists(rep):
    det = subprocess.run(["ncu", "-i", rep, "--page", "details"], capture_output=True, text=True).stdout
    summ = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep, "25"], capture_output=True, text=True).stdout
    open(os.path.join(P, f"{tag}_ncu_full_conv_umma_k11.txt"), "w").write(
        "ncu --set full --clock-control none, conv_umma_kernel, stage-2 shape (B=32, L=48001, C=128, k=11, snake + stats)\n\n" + summ + "\n" + det)
print("launch shares:"); print(open(os.path.join(P, f"{tag}_launch_shares_c2.csv")).read()[:700])
print(open(os.path.join(P, "conv_umma_traffic.json")).read())

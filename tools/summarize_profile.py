"""Turn ncu captures (gpurun_out/*.ncu-rep, launch-list csv) into the committed summaries under profiles/.

    python tools/summarize_profile.py <k2.ncu-rep> <launches.csv> <rollouts per K2 launch> [tag]

Writes profiles/<tag>_k2_raw.csv (selected raw metrics), profiles/<tag>_launches.csv (kernel, launches,
total / mean device time, share) and profiles/<tag>_summary.json (read by bench.py for roofline.traffic).
"""
import csv, json, os, subprocess, sys, collections

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, launches, per_launch = sys.argv[1], sys.argv[2], int(sys.argv[3])
tag = sys.argv[4] if len(sys.argv) > 4 else "r01"
out = os.path.join(ROOT, "profiles")
os.makedirs(out, exist_ok=True)

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
keep = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma", "smsp__issue_active.avg.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled", "sm__throughput.avg.pct", "lts__t_bytes.sum", "sm__cycles_elapsed.max",
        "launch__shared_mem_per_block_dynamic", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
sel = [(h, u, v) for h, u, v in zip(hdr, units, vals) if any(h.startswith(k) for k in keep) and "per_second" not in h]
with open(os.path.join(out, f"{tag}_k2_raw.csv"), "w") as f:
    w = csv.writer(f)
    w.writerow(("metric", "unit", "value"))
    w.writerows(sel)
d = {h: (u, v) for h, u, v in sel}


def num(name):
    u, v = d[name]
    x = float(v.replace(",", ""))
    return x * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "s": 1, "ns": 1e-9}.get(u, 1)


traffic = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
summary = {"k2_kernel": d["Kernel Name"][1] if "Kernel Name" in d else None,
           "k2_rollouts_per_launch": per_launch, "k2_dram_bytes_per_launch": traffic,
           "k2_dram_bytes_per_rollout": traffic / per_launch, "k2_duration_s_under_ncu": num("gpu__time_duration.sum"),
           "source": os.path.basename(rep)}

agg = collections.OrderedDict()
with open(launches) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    k = r["Kernel Name"]
    v = float(r["Metric Value"].replace(",", "")) * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}[r["Metric Unit"]]
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values()) or 1.0
with open(os.path.join(out, f"{tag}_launches.csv"), "w") as f:
    w = csv.writer(f)
    w.writerow(("kernel", "launches", "total_ms", "mean_ms", "share"))
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        w.writerow((k[:110], n, f"{t * 1e3:.4f}", f"{t / n * 1e3:.4f}", f"{t / tot:.4f}"))
k2 = [t for k, (n, t) in agg.items() if "am_rollout" in k]
summary["k2_share_of_launch_list"] = sum(k2) / tot if k2 else None
with open(os.path.join(out, f"{tag}_summary.json"), "w") as f:
    json.dump(summary, f, indent=1)
print(json.dumps(summary, indent=1))

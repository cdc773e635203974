"""One ncu --set full capture (.ncu-rep) -> profiles/<tag>_raw.csv with the metrics the roofline discussion uses,
and a one-line JSON summary on stdout.     python tools/summarize_side.py <rep> <tag> <units per launch> <algorithmic bytes per unit>
"""
import csv, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, tag, units, algo = sys.argv[1], sys.argv[2], float(sys.argv[3]), float(sys.argv[4])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, unit, val = rows[0], rows[1], rows[2]
keep = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit", "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma",
        "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "smsp__average_warps_issue_stalled",
        "smsp__average_warp_latency_issue_stalled", "smsp__average_warp_pct_issue_stalled",
        "sm__throughput.avg.pct", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.max",
        "launch__shared_mem_per_block", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
        "smsp__sass_average_data_bytes_per_sector_mem_global")
sel = [(h, u, v) for h, u, v in zip(hdr, unit, val) if any(h.startswith(k) for k in keep) and "per_second" not in h]
with open(os.path.join(ROOT, "profiles", f"{tag}_raw.csv"), "w") as f:
    w = csv.writer(f)
    w.writerow(("metric", "unit", "value"))
    w.writerows(sel)
d = {h: (u, v) for h, u, v in sel}
SC = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "s": 1, "ns": 1e-9, "msecond": 1e-3,
      "usecond": 1e-6, "nsecond": 1e-9, "second": 1}


def num(name):
    u, v = d[name]
    return float(v.replace(",", "")) * SC.get(u, 1)


def opt(name):
    try:
        return float(d[name][1].replace(",", ""))
    except Exception:  # noqa: BLE001
        return None


traffic, dur = num("dram__bytes_read.sum") + num("dram__bytes_write.sum"), num("gpu__time_duration.sum")
print(json.dumps({"tag": tag, "kernel": d["Kernel Name"][1][:90], "units_per_launch": units, "duration_s_under_ncu": dur,
                  "dram_bytes_per_launch": traffic, "dram_bytes_per_unit": traffic / units, "algorithmic_bytes_per_unit": algo,
                  "traffic_over_algorithmic": traffic / units / algo, "dram_GBps_under_ncu": traffic / dur / 1e9,
                  "dram_pct_of_peak": opt("dram__throughput.avg.pct_of_peak_sustained_elapsed"),
                  "warps_active_pct": opt("sm__warps_active.avg.pct_of_peak_sustained_active"),
                  "issue_active_pct": opt("smsp__issue_active.avg.pct"),
                  "tensor_pipe_pct": opt("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
                  "registers": opt("launch__registers_per_thread"), "grid": opt("launch__grid_size"),
                  "block": opt("launch__block_size"), "source": os.path.basename(rep)}))

#!/usr/bin/env python3
"""tools/summarise_profile.py <round tag, e.g. r01g> — turns gpurun_out/prof_<tag>.ncu-rep, launches_<tag>.csv and bench_<tag>.json into the
committed files under profiles/ (selected --page raw metrics, launch shares, DRAM traffic per traceKernel launch)."""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
raw = subprocess.run(["ncu", "-i", os.path.join(G, f"prof_{tag}.ncu-rep"), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = rows[0]
cols = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_issued.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "sm__cycles_active.avg", "sm__cycles_elapsed.max"]
idx = [h.index(c) for c in cols]
out = [[r[i] for i in idx] for r in rows]
summary = os.path.join(P, f"{tag}_trace_shade_ncu_full_summary.csv")
with open(summary, "w", newline="") as f:
    csv.writer(f).writerows(out)
for r in out[2:]:
    print(r[0][:24], r[3:])
units = out[1]
mul = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}
tr = [r for r in out[2:] if "traceKernel" in r[0]]
traffic = sum(float(r[4]) * mul[units[4]] + float(r[5]) * mul[units[5]] for r in tr) / len(tr)
def avg(col):
    return sum(float(r[cols.index(col)]) for r in tr) / len(tr)
json.dump({"kernel": "traceKernel", "dram_bytes_per_launch": traffic, "launches_captured": len(tr),
           "ipc_per_sm": avg("sm__inst_issued.avg.per_cycle_active"), "active_threads_per_warp_inst": avg("smsp__thread_inst_executed_per_inst_executed.ratio"),
           "l1_wavefront_pct": avg("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"), "l1_hit_pct": avg("l1tex__t_sector_hit_rate.pct"),
           "l2_hit_pct": avg("lts__t_sector_hit_rate.pct"), "dram_pct_of_peak": avg("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
           "source": f"profiles/{tag}_trace_shade_ncu_full_summary.csv (ncu --set full --clock-control none, bounces 2-3 of a frame; dram__bytes_read.sum + dram__bytes_write.sum)"},
          open(os.path.join(P, "traffic.json"), "w"), indent=1)
print("traffic per traceKernel launch:", traffic)
shutil.copy(os.path.join(G, f"bench_{tag}.json"), os.path.join(P, f"{tag}_bench.json"))
shutil.copy(os.path.join(G, f"launches_{tag}.csv"), os.path.join(P, f"{tag}_launches.csv"))
d = json.load(open(os.path.join(P, f"{tag}_bench.json")))
print("bench:", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], "Mrays/s", d["mrays_per_s"], "roofline", d["roofline"]["achieved"], d["roofline"]["frac"])
print("stages:", d["stage_ms_per_step"], "cpu:", d.get("cpu_baseline", {}).get("value"))
rows = list(csv.reader(open(os.path.join(P, f"{tag}_launches.csv"))))
st = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hh = rows[st]
kn, mv, mu = hh.index("Kernel Name"), hh.index("Metric Value"), hh.index("Metric Unit")
agg = collections.Counter()
for r in rows[st + 1:]:
    if len(r) > mv:
        agg[r[kn].split("(")[0].replace("void ", "")] += float(r[mv].replace(",", "")) * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(r[mu], 1)
tot = sum(agg.values())
print("launch shares:", [(k, round(100 * v / tot, 1)) for k, v in agg.most_common(7)])

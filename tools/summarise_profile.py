#!/usr/bin/env python3
"""tools/summarise_profile.py <tag> [c5-tag] — turns gpurun_out/prof_<tag>.ncu-rep (one whole frame of the bench workload under `ncu --set full`),
launches_<tag>.csv and bench_<tag>.json into the committed files under profiles/: selected --page raw metrics per launch, launch shares, and
profiles/traffic.json = per-kernel averages (DRAM bytes, L2 bytes, duration, issue / occupancy figures) that bench.py turns into dram_frac / l2_frac.
With a second tag, gpurun_out/prof_<c5-tag>.ncu-rep (traceKernel launches of the 20 M-triangle 4K config) is summarised under "c5_traceKernel"."""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
COLS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors.sum",
        "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_issued.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "sm__cycles_active.avg", "sm__cycles_elapsed.max"]
MUL = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1, "Tbyte": 1e12, "us": 1.0, "ns": 1e-3, "ms": 1e3, "s": 1e6, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}


def raw_rows(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h = rows[0]
    cols = [c for c in COLS if c in h]
    idx = [h.index(c) for c in cols]
    return cols, [[r[i] for i in idx] for r in rows]


def per_kernel(cols, out, hbm_peak):
    units = dict(zip(cols, out[1]))

    def val(r, c):
        if c not in cols or r[cols.index(c)] in ("", "n/a"):
            return None
        return float(r[cols.index(c)].replace(",", "")) * MUL.get(units[c], 1)

    groups = collections.OrderedDict()
    for r in out[2:]:
        name = r[0].split("(")[0].replace("void ", "").split("<")[0]
        groups.setdefault(name, []).append(r)
    res = {}
    for name, rs in groups.items():
        def avg(c):
            v = [val(r, c) for r in rs]
            v = [x for x in v if x is not None]
            return sum(v) / len(v) if v else None
        dur = avg("gpu__time_duration.sum")
        rd, wr = avg("dram__bytes_read.sum"), avg("dram__bytes_write.sum")
        dram = (rd or 0) + (wr or 0) if rd is not None or wr is not None else None
        res[name] = {"launches_captured": len(rs), "duration_us": dur, "dram_bytes_per_launch": dram, "lts_bytes_per_launch": (32.0 * avg("lts__t_sectors.sum")) if avg("lts__t_sectors.sum") is not None else None,
                     "dram_frac_of_peak": (dram / (dur * 1e-6) / 1e9 / hbm_peak) if dram is not None and dur else None,
                     "ipc_per_sm": avg("sm__inst_issued.avg.per_cycle_active"),
                     "active_threads_per_warp_inst": avg("smsp__thread_inst_executed_per_inst_executed.ratio"),
                     "warps_active_pct": avg("sm__warps_active.avg.pct_of_peak_sustained_active"),
                     "long_scoreboard_per_issue": avg("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
                     "l1_wavefront_pct": avg("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"), "l1_hit_pct": avg("l1tex__t_sector_hit_rate.pct"),
                     "l2_hit_pct": avg("lts__t_sector_hit_rate.pct"), "dram_pct_of_peak": avg("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
                     "registers": avg("launch__registers_per_thread")}
    return res


def main():
    tag = sys.argv[1]
    c5tag = sys.argv[2] if len(sys.argv) > 2 else None
    try:
        hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        hbm_peak = 7700.0
    cols, out = raw_rows(os.path.join(G, f"prof_{tag}.ncu-rep"))
    with open(os.path.join(P, f"{tag}_frame_ncu_full_summary.csv"), "w", newline="") as f:
        csv.writer(f).writerows(out)
    kernels = per_kernel(cols, out, hbm_peak)
    tj = {"source": f"profiles/{tag}_frame_ncu_full_summary.csv (ncu --set full --clock-control none: every traceKernel / shadeKernel / gatherKernel / finalizeKernel "
                    "launch of one frame of the bench workload; per-kernel averages; dram = dram__bytes_read.sum + dram__bytes_write.sum, l2 = 32 B x lts__t_sectors.sum)",
          "hbm_peak_gbs": hbm_peak, "kernels": kernels}
    if c5tag:
        c5cols, c5out = raw_rows(os.path.join(G, f"prof_{c5tag}.ncu-rep"))
        with open(os.path.join(P, f"{c5tag}_ncu_full_summary.csv"), "w", newline="") as f:
            csv.writer(f).writerows(c5out)
        k5 = per_kernel(c5cols, c5out, hbm_peak).get("traceKernel")
        if k5:
            k5["source"] = f"profiles/{c5tag}_ncu_full_summary.csv (traceKernel launches of the 20 M-triangle 3840x2160 config)"
            tj["c5_traceKernel"] = k5
    else:
        try:
            old = json.load(open(os.path.join(P, "traffic.json")))
            if "c5_traceKernel" in old:
                tj["c5_traceKernel"] = old["c5_traceKernel"]
        except Exception:
            pass
    json.dump(tj, open(os.path.join(P, "traffic.json"), "w"), indent=1)
    for k, v in kernels.items():
        print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items()})
    for name in (f"bench_{tag}.json", f"launches_{tag}.csv"):
        if os.path.exists(os.path.join(G, name)):
            shutil.copy(os.path.join(G, name), os.path.join(P, f"{tag}_" + name.replace(f"_{tag}", "")))
    lp = os.path.join(P, f"{tag}_launches.csv")
    if os.path.exists(lp):
        rows = list(csv.reader(open(lp)))
        st = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
        hh = rows[st]
        kn, mv, mu = hh.index("Kernel Name"), hh.index("Metric Value"), hh.index("Metric Unit")
        agg = collections.Counter()
        for r in rows[st + 1:]:
            if len(r) > mv:
                agg[r[kn].split("(")[0].replace("void ", "")] += float(r[mv].replace(",", "")) * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(r[mu], 1)
        tot = sum(agg.values())
        print("launch shares:", [(k, round(100 * v / tot, 1)) for k, v in agg.most_common(8)])


if __name__ == "__main__":
    main()

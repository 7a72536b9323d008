#!/usr/bin/env python
"""Per-source-line totals from an ncu report's source page:
   ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv ; python tools/source_hotspots.py src.csv [kernel-substring] [top]
Prints, for each CUDA source line, warp instructions executed, thread instructions, active threads per instruction and stall samples."""
import csv
import sys


KEY = int(__import__("os").environ.get("HOT_KEY", 2))


def main():
    path, want, top = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else ""), int(sys.argv[3]) if len(sys.argv) > 3 else 40
    rows = list(csv.reader(open(path)))
    kernel, hdr, out = None, None, {}
    for r in rows:
        if r and r[0] == "Function Name":
            kernel = r[1]
            continue
        if r and r[0] == "Line No":
            hdr = r
            continue
        if not hdr or kernel is None or want not in kernel or len(r) < len(hdr) or not r[0]:
            continue
        n = len(hdr) - 4                               # per-line aggregate row (columns 0-3: line, source, address, sass); the source
        d = dict(zip(hdr[4:], r[-n:]))                 # text is not escaped (asm strings hold quotes), so the metrics are taken from the right
        if not r[0].isdigit() or not d["Instructions Executed"].isdigit():
            continue
        out.setdefault(kernel, []).append((int(r[0]), r[1], int(d["Instructions Executed"]), int(d["Thread Instructions Executed"]),
                                           int(d["# Samples"]), int(d.get("stall_long_sb", 0) or 0)))
    for k, lines in out.items():
        ti, tt, ts = sum(x[2] for x in lines), sum(x[3] for x in lines), sum(x[4] for x in lines)
        print(f"== {k[:60]}: {ti} warp inst, {tt / max(1, ti):.1f} threads/inst, {ts} samples")
        for ln, src, wi, th, smp, lsb in sorted(lines, key=lambda x: -x[KEY])[:top]:
            print(f"{ln:5d} inst {100 * wi / ti:5.1f}%  thr/inst {th / max(1, wi):5.1f}  samples {100 * smp / max(1, ts):5.1f}%  long_sb {100 * lsb / max(1, ts):4.1f}%  | {src.strip()[:110]}")


if __name__ == "__main__":
    main()

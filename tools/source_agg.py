#!/usr/bin/env python
"""Per-source-line totals of one kernel SUMMED over all of its captured launches, with the two leading stall reasons of each line:
   ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv ; python tools/source_agg.py src.csv <kernel-substring> [top] [inst|samples]"""
import collections
import csv
import sys


def main():
    path, want = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    key = 0 if len(sys.argv) > 4 and sys.argv[4] == "inst" else 2
    rows = list(csv.reader(open(path)))
    kernel, hdr = None, None
    agg = collections.defaultdict(lambda: [0, 0, 0, collections.Counter()])
    for r in rows:
        if r and r[0] == "Function Name":
            kernel = r[1]
            continue
        if r and r[0] == "Line No":
            hdr = r
            continue
        if not hdr or not kernel or want not in kernel or len(r) < len(hdr) or not r[0].isdigit():
            continue
        n = len(hdr) - 4
        d = dict(zip(hdr[4:], r[-n:]))
        if not d["Instructions Executed"].isdigit():
            continue
        a = agg[(int(r[0]), r[1].strip()[:100])]
        a[0] += int(d["Instructions Executed"])
        a[1] += int(d["Thread Instructions Executed"])
        a[2] += int(d["# Samples"])
        for k, v in d.items():
            if k.startswith("stall_") and v.isdigit():
                a[3][k] += int(v)
    ti, ts = sum(a[0] for a in agg.values()), sum(a[2] for a in agg.values())
    tot = collections.Counter()
    for a in agg.values():
        tot.update(a[3])
    print(f"{want}: {ti} warp instructions, {sum(a[1] for a in agg.values()) / max(1, ti):.1f} threads/inst, {ts} samples; stall mix (% of samples):",
          [(k[6:], round(100 * v / max(1, ts), 1)) for k, v in tot.most_common(8)])
    for (ln, src), a in sorted(agg.items(), key=lambda x: -x[1][key])[:top]:
        t = a[3].most_common(2) + [("stall_-", 0)] * 2
        print(f"{ln:4d} inst {100 * a[0] / ti:4.1f}% thr {a[1] / max(1, a[0]):4.1f} samples {100 * a[2] / max(1, ts):4.1f}% "
              f"{t[0][0][6:]}={100 * t[0][1] / max(1, ts):3.1f} {t[1][0][6:]}={100 * t[1][1] / max(1, ts):3.1f} | {src}")


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel over ONE step.

  python profiles/summarize_launches.py gpurun_out/launches.csv 488 > profiles/rNN_launches_summary.txt

The last `launches_per_step` rows of the list form one full train step (any window of that length over
the steady-state eager loop contains every launch of the step exactly once)."""
import collections
import csv
import re
import sys


def main():
    path, per_step = sys.argv[1], int(sys.argv[2])
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    h = rows[hdr]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    data = []
    for r in rows[hdr + 1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] in ("ns", "nsecond") else v
        data.append((re.sub(r"\(.*", "", r[ki]), v))
    step = data[-per_step:]
    agg = collections.defaultdict(lambda: [0.0, 0])
    for k, v in step:
        agg[k][0] += v
        agg[k][1] += 1
    tot = sum(v for _, v in step)
    print(f"one full train step = {len(step)} launches, {tot / 1e3:.2f} ms summed kernel time (cold-cache, serialised)")
    for k, (t, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{t:10.1f} us {100 * t / tot:5.1f}% x{n:4d} avg {t / n:8.1f} us  {k}")


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: one forward step (im2col .. next im2col),
aggregated per kernel, plus the per-launch list.  Usage: launch_summary.py launches.csv [--list]"""
import collections
import csv
import re
import sys


def load(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            v = float(row["Metric Value"].replace(",", ""))
            u = row["Metric Unit"]
            v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
            name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("<unnamed>::", "").replace("void ", "")
            rows.append((name, row["Grid Size"], v))
    return rows


def main():
    rows = load(sys.argv[1])
    idx = [i for i, x in enumerate(rows) if "im2col" in x[0]]
    step = rows[idx[0]:idx[1]]
    tot = sum(x[2] for x in step)
    print(f"one forward step: {len(step)} launches, {tot / 1e3:.2f} ms summed kernel time (ncu: cold cache, serialised)")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for x in step:
        agg[x[0]][0] += 1
        agg[x[0]][1] += x[2]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1]:10.1f} us {100 * v[1] / tot:5.1f}%  n={v[0]:3d}  {k[:100]}")
    if "--list" in sys.argv:
        for i, x in enumerate(step):
            print(f"{i:4d} {x[2]:9.1f} us {x[1]:>16s} {x[0][:80]}")


if __name__ == "__main__":
    main()

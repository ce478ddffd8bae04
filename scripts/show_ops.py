#!/usr/bin/env python
"""Print a range of per-op timings from gpurun_out/op_times_situ.json: show_ops.py lo hi [lo hi ...]"""
import json
import sys

r = json.load(open("gpurun_out/op_times_situ.json"))
b = [int(x) for x in sys.argv[1:]] or [0, len(r)]
for lo, hi in zip(b[::2], b[1::2]):
    for i, x in enumerate(r):
        if lo <= i <= hi:
            print(i, x["op"][:14], x.get("M"), x.get("N"), x.get("K"), x.get("taps"), x.get("act"), x.get("map"), round(x["us"], 1))

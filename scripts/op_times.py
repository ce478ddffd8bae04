#!/usr/bin/env python
"""Per-op device times of one forward step (CUDA events, warm, median of R repeats): which ops of the plan sit how far
from their tensor-pipe / HBM floor.  Run on the GPU box: python scripts/op_times.py [--encoder vitl] [--frames 32]."""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from video_depth_anything_b200 import lib as L  # noqa: E402
from video_depth_anything_b200.engine import ForwardPlan, PackedWeights  # noqa: E402
from bench import build_model  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--encoder", default="vitl")
    ap.add_argument("--frames", type=int, default=32)
    ap.add_argument("--size", type=int, default=518)
    ap.add_argument("--width", type=int, default=0, help="network width if not square (924 = configs[3]'s 518x924)")
    ap.add_argument("--repeats", type=int, default=7)
    ap.add_argument("--out", default="gpurun_out/op_times.json")
    ap.add_argument("--in-situ", action="store_true", help="time the ops inside back-to-back steps (events between ops, no flush)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    model = build_model(a.encoder, dev)  # the shipped configuration of that encoder (features, out_channels)
    pw = PackedWeights(model, dev, bf16=True)
    plan = ForwardPlan(pw, 1, a.frames, a.size, a.width or a.size, dev)
    plan.x_in.normal_()
    lib = plan._lib
    st = L.stream_ptr()
    for _ in range(2):
        plan.launch()
    torch.cuda.synchronize()
    rows = []
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    situ = None
    if a.in_situ:
        n = len(plan.ops)
        acc = [[] for _ in range(n)]
        for _ in range(a.repeats):
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
            evs[0].record()
            for idx, (fn, args) in enumerate(plan.ops):
                rc = fn(C.byref(args[0]), st) if fn is lib.vda_gemm else fn(*args, st)
                L.check(rc, "op %d" % idx)
                evs[idx + 1].record()
            torch.cuda.synchronize()
            for idx in range(n):
                acc[idx].append(evs[idx].elapsed_time(evs[idx + 1]) * 1e3)
        situ = [sorted(x)[len(x) // 2] for x in acc]
    for idx, (fn, args) in enumerate(plan.ops):
        ts = []
        for _ in range(0 if situ else a.repeats):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(C.byref(args[0]), st) if fn is lib.vda_gemm else fn(*args, st)
            e1.record()
            L.check(rc, "op %d" % idx)
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        us = situ[idx] if situ else ts[len(ts) // 2]
        if fn is lib.vda_gemm:
            d = args[0]
            fl = 2.0 * d.M * d.N * d.K * d.taps * d.ksplit
            rows.append(dict(i=idx, op="gemm", M=d.M, N=d.N, K=d.K, taps=d.taps, act=d.act, map=d.out_map, us=us,
                             flops=fl))
        else:
            rows.append(dict(i=idx, op=getattr(fn, "__name__", str(fn)), us=us))
    total = sum(r["us"] for r in rows)
    # aggregate identical shapes
    agg = {}
    for r in rows:
        key = (r["op"], r.get("M"), r.get("N"), r.get("K"), r.get("taps"), r.get("act"), r.get("map"))
        g = agg.setdefault(key, dict(n=0, us=0.0, fl=0.0))
        g["fl"] += r.get("flops", 0.0)
        g["n"] += 1
        g["us"] += r["us"]
    print(f"{len(rows)} ops, {total / 1e3:.2f} ms summed ({'in situ: events between the ops of back-to-back steps' if situ else 'L2 flushed before each op'})")
    for key, g in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
        op, M, N, K, taps, act, mp = key
        shape = f"M={M} N={N} K={K} taps={taps} act={act} map={mp}" if op == "gemm" else ""
        tf = f"{g['fl'] / g['us'] / 1e6:7.0f} TF/s" if op == "gemm" else ""
        print(f"{g['us']:9.1f} us {100 * g['us'] / total:5.1f}%  n={g['n']:3d} {g['us'] / g['n']:8.1f} us/op {tf}  {op} {shape}")
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(rows, open(a.out, "w"))


if __name__ == "__main__":
    main()

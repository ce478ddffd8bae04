#!/usr/bin/env python
"""Isolated timing of the tcgen05 GEMM on the encoder shapes (CUDA events, L2 flushed between launches by working-set size)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from video_depth_anything_b200 import lib as L  # noqa: E402

if os.environ.get("VDA_LIB"):   # A/B against an experimental build of the library
    L.LIB_PATH = os.environ["VDA_LIB"]
dev = "cuda"
L.require_device(dev)
lib = L.load()
M = 43840
shapes = [("qkv", 3072, 1024, "store"), ("proj", 1024, 1024, "resid"), ("fc1", 4096, 1024, "gelu"), ("fc2", 1024, 4096, "resid"),
          ("projS", 1024, 1024, "store"), ("fc2S", 1024, 4096, "store"), ("projRelu", 1024, 1024, "relu2"),
          ("projRes", 1024, 1024, "res"), ("projResIn", 1024, 1024, "resin"), ("mmRes", 256, 256, "res"), ("mmS", 256, 256, "store"), ("fc1S", 4096, 1024, "store")]
if len(sys.argv) > 1:
    shapes = [s for s in shapes if s[0] in sys.argv[1:]]
for name, N, K, kind in shapes:
    A = torch.randn(M, K, device=dev).bfloat16()
    B = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device=dev)
    d = L.GemmDesc()
    d.A, d.B, d.lda, d.ldb, d.rowsA = A.data_ptr(), B.data_ptr(), K, K, M
    d.a_dtype = d.b_dtype = L.BF16
    d.M, d.N, d.K, d.taps, d.ksplit = M, N, K, 1, 1
    d.bias = bias.data_ptr()
    d.round_bf16, d.out_dtype, d.use_tc = 1, L.BF16, 1
    keep = []
    if kind == "resid":
        resid = torch.randn(M, N, device=dev)
        gam = torch.rand(N, device=dev)
        d.resid, d.gamma, d.ldresid = resid.data_ptr(), gam.data_ptr(), N
        keep += [resid, gam]
    else:
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        d.out, d.ldo = out.data_ptr(), N
        d.act = L.ACT_GELU if kind == "gelu" else L.ACT_NONE
        keep.append(out)
        if kind == "relu2":
            o2 = torch.empty_like(out)
            d.out_relu = o2.data_ptr()
            keep.append(o2)
        if kind in ("res", "resin"):
            r = out if kind == "resin" else torch.randn(M, N, device=dev).bfloat16()
            d.res, d.ldres = r.data_ptr(), N
            keep.append(r)
    for _ in range(3):
        L.gemm(d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = int(os.environ.get("GEMM_BENCH_ITERS", "20"))
    e0.record()
    for _ in range(n):
        L.gemm(d)
    e1.record()
    clk = pw = ""
    if n >= 200:  # long enough to sample the SM clock and board power under load (the step runs at the power cap)
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(0)
        cs, ps = [], []
        while not e1.query():
            cs.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
            ps.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
        if cs:
            cs, ps = sorted(cs[len(cs) // 3:]), sorted(ps[len(ps) // 3:])
            clk, pw = f"  sm {cs[len(cs) // 2]} MHz", f"  {ps[len(ps) // 2]:.0f} W"
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    print(f"{name:5s} M={M} N={N} K={K} {kind:6s}: {us:8.1f} us  {2.0 * M * N * K / us / 1e6:8.1f} TFLOP/s{clk}{pw}  (VDA_NO_PAIR={os.environ.get('VDA_NO_PAIR', '0')})")

#!/usr/bin/env python
"""K / N sweep of the tcgen05 GEMM (plain store epilogue) to separate mainloop rate from per-tile overhead."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from video_depth_anything_b200 import lib as L  # noqa: E402

if os.environ.get("VDA_LIB"):   # A/B against an experimental build of the library
    L.LIB_PATH = os.environ["VDA_LIB"]
dev = "cuda"
L.require_device(dev)
M = 43840
for N, K in [(1024, 1024), (1024, 4096), (1024, 16384), (2048, 8192), (4096, 1024), (4096, 4096)]:
    A = torch.randn(M, K, device=dev).bfloat16()
    B = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    d = L.GemmDesc()
    d.A, d.B, d.lda, d.ldb, d.rowsA = A.data_ptr(), B.data_ptr(), K, K, M
    d.a_dtype = d.b_dtype = L.BF16
    d.M, d.N, d.K, d.taps, d.ksplit = M, N, K, 1, 1
    d.round_bf16, d.out_dtype, d.use_tc = 1, L.BF16, 1
    d.out, d.ldo = out.data_ptr(), N
    for _ in range(2):
        L.gemm(d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 30
    e0.record()
    for _ in range(n):
        L.gemm(d)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    # cuBLAS on the same shape for reference
    torch.matmul(A, B.T)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        torch.matmul(A, B.T)
    e1.record()
    torch.cuda.synchronize()
    us2 = e0.elapsed_time(e1) * 1e3 / n
    fl = 2.0 * M * N * K / 1e6
    print(f"N={N:5d} K={K:6d}: ours {us:8.1f} us {fl / us:7.1f} TFLOP/s | cuBLAS {us2:8.1f} us {fl / us2:7.1f} TFLOP/s  (NO_PAIR={os.environ.get('VDA_NO_PAIR', '0')})")

#!/usr/bin/env python
"""Epilogue timeline of the pair GEMM (needs a library built with -DTC2_TRACE; debugging aid): per tile of CTA 0, how long
epilogue warp 0 waited for the accumulator and how long its share of the epilogue took.
VDA_LIB=scripts/micro/libvda_tc2trace.so python scripts/gemm_trace.py [gelu|store|resid|store4096]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from video_depth_anything_b200 import lib as L  # noqa: E402

if os.environ.get("VDA_LIB"):
    L.LIB_PATH = os.environ["VDA_LIB"]
lib = L.load()
dev = "cuda"
kind = sys.argv[1] if len(sys.argv) > 1 else "gelu"
deep = kind in ("resid", "store4096")   # fc2's shape (N = 1024, K = 4096) instead of fc1's
M, N, K = 43840, (1024 if deep else 4096), (4096 if deep else 1024)
A = torch.randn(M, K, device=dev).bfloat16()
B = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
bias = torch.randn(N, device=dev)
d = L.GemmDesc()
d.A, d.B, d.lda, d.ldb, d.rowsA = A.data_ptr(), B.data_ptr(), K, K, M
d.a_dtype = d.b_dtype = L.BF16
d.M, d.N, d.K, d.taps, d.ksplit = M, N, K, 1, 1
d.bias = bias.data_ptr()
d.round_bf16, d.out_dtype, d.use_tc = 1, L.BF16, 1
if kind == "resid":
    resid = torch.randn(M, N, device=dev)
    gam = torch.rand(N, device=dev)
    d.resid, d.gamma, d.ldresid = resid.data_ptr(), gam.data_ptr(), N
else:
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    d.out, d.ldo = out.data_ptr(), N
    d.act = L.ACT_GELU if kind == "gelu" else L.ACT_NONE
for _ in range(3):
    L.gemm(d)
torch.cuda.synchronize()
buf = (C.c_longlong * (64 * 4))()
print("rc", C.CDLL(lib._name).vda_debug_tc2_trace(buf))
rows = [[buf[t * 4 + e] for e in range(3)] for t in range(64)]
rows = [r for r in rows if r[0] and r[2] > r[0]]
print(f"{kind}: {len(rows)} tiles of CTA 0")
for t, r in enumerate(rows[:40]):
    per = (rows[t + 1][0] - r[0]) if t + 1 < len(rows) else 0
    print(f"tile {t:2d}: waited for the accumulator {r[1] - r[0]:6d} clk, epilogue {r[2] - r[1]:6d} clk, tile period {per:6d} clk")

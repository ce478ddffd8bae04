#!/usr/bin/env python
"""Phase timeline of the flash-attention softmax warps (needs a build with -DFA_TRACE; debugging aid, not product)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from video_depth_anything_b200 import lib as L  # noqa: E402

if os.environ.get("VDA_LIB"):
    L.LIB_PATH = os.environ["VDA_LIB"]
lib = L.load()
dev = torch.device("cuda:0")
frames, ntok, heads = 32, 1370, 16
qkv = (torch.randn(frames * ntok, 3 * heads * 64, device=dev) * 1.3).bfloat16()
o = torch.empty(frames * ntok, heads * 64, device=dev, dtype=torch.bfloat16)
for _ in range(3):
    L.check(lib.vda_attention(qkv.data_ptr(), o.data_ptr(), L.BF16, frames, ntok, heads, 1, 1, L.stream_ptr()), "attn")
torch.cuda.synchronize()
buf = (C.c_longlong * (4 * 16 * 8))()
NSTREAM = 2  # streams per CTA (key halves)
raw = C.CDLL(lib._name)
print("rc", raw.vda_debug_fa_trace(buf))
names = ["loop", "S ready", "read+pack", "max", "exp", "PVdone", "store+arrive"]
t0 = min(buf[(w * 16 + 2) * 8] for w in range(NSTREAM))
for j in range(2, 9):
    for w in range(NSTREAM):
        ev = [buf[(w * 16 + j) * 8 + e] for e in range(7)]
        d = [ev[i + 1] - ev[i] for i in range(6)]
        print(f"j={j} stream{w} start={ev[0] - t0:6d}  " + "  ".join(f"{n}={x:5d}" for n, x in zip(names[1:], d)) +
              f"  total={ev[6] - ev[0]:5d}")
    print()

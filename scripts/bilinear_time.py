"""Isolated timing of the largest fusion-block upsample (148 -> 296, 256 channels, 32 frames, padded in / dense out)."""
import sys, torch
sys.path.insert(0, "/root/repo")
from video_depth_anything_b200 import lib as L
lib = L.load(); dev = torch.device("cuda:0")
for (frames, ih, iw, oh, ow, C) in ((32, 148, 148, 296, 296, 256), (32, 74, 74, 148, 148, 256), (32, 148, 148, 296, 296, 64)):
    x = torch.randn(frames * (ih + 2) * (iw + 2), C, device=dev).bfloat16()
    y = torch.empty(frames * oh * ow, C, device=dev, dtype=torch.bfloat16)
    f = lambda: L.check(lib.vda_bilinear(x.data_ptr(), L.BF16, y.data_ptr(), L.BF16, frames, ih, iw, 1, oh, ow, 0, C, L.stream_ptr()), "bilinear")
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): f()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    gb = (x.numel() + y.numel()) * 2 / 1e9
    print(f"bilinear {ih}->{oh} C={C}: {us:.1f} us, {gb / us * 1e6:.0f} GB/s (in + out)")

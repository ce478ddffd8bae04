#!/usr/bin/env python
"""Eager launch loop vs CUDA-graph replay of one forward plan (device time per clip, CUDA events)."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from video_depth_anything_b200.engine import ForwardPlan, PackedWeights  # noqa: E402
from video_depth_anything_b200.video_depth import VideoDepthAnything  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
enc = sys.argv[1] if len(sys.argv) > 1 else "vitl"
model = VideoDepthAnything(encoder=enc).to(dev)
plan = ForwardPlan(PackedWeights(model, dev, bf16=True), 1, 32, 518, 518, dev)
plan.x_in.normal_()


def timed(fn, n=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print("eager  %.3f ms/clip" % timed(plan.launch))
ref = plan.depth.clone()
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    plan.launch()
torch.cuda.current_stream().wait_stream(s)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    plan.launch()
print("graph  %.3f ms/clip" % timed(g.replay))
print("eager2 %.3f ms/clip" % timed(plan.launch))
print("graph output identical:", bool(torch.equal(ref, plan.depth)))

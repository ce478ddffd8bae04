#!/usr/bin/env python
"""Where a streaming step's time goes: host enqueue time vs device time of `StreamingDepthDriver.run` (results stay on
the device, so the only host waits are the ones inside the loop)."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import build_model
from scripts.driver_bench import synth
from video_depth_anything_b200.stream import StreamingDepthDriver

enc = sys.argv[1] if len(sys.argv) > 1 else "vitl"
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
model = build_model(enc, dev)
frames = synth(160, 518, 518)
drv = StreamingDepthDriver(model, dev)
kw = dict(inference_length=32, keyframe_list=[0, 12], align_each_new_frame=False, return_device=True)
drv.run(frames[:48], **kw); torch.cuda.synchronize()
for n in (48, 160):
    t0 = time.perf_counter()
    out = drv.run(frames[:n], **kw)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"{enc} n={n}: host enqueue {1e3 * (t1 - t0):.1f} ms, total {1e3 * (t2 - t0):.1f} ms")

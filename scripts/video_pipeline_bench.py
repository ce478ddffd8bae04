#!/usr/bin/env python
"""File -> frames -> depth -> pictures -> file on one GPU: stage times of
`read_video_frames` (host decode into page-locked memory), `infer_video_depth` on page-locked vs pageable frames,
`depth_pictures` (device visualisation) and the host encoder, plus the two visualisation kernels alone against the HBM
roofline (CUDA events; 4 B read per pixel for the min/max pass, 4 B + 1 or 3 B for the picture pass)."""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--encoder", default="vits")
    ap.add_argument("--frames", type=int, default=256)
    ap.add_argument("--h", type=int, default=720)
    ap.add_argument("--w", type=int, default=1280)
    a = ap.parse_args()
    import cv2
    from bench import build_model
    from scripts.driver_bench import synth
    from video_depth_anything_b200 import lib as L
    from video_depth_anything_b200 import video_io as vio
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    model = build_model(a.encoder, dev)
    tmp = tempfile.mkdtemp()
    src = os.path.join(tmp, "src.mp4")
    wr = cv2.VideoWriter(src, cv2.VideoWriter_fourcc(*"mp4v"), 30.0, (a.w, a.h))
    for f in synth(a.frames, a.h, a.w):
        wr.write(f)
    wr.release()
    res = {"config": f"{a.encoder}, {a.frames} frames {a.h}x{a.w} mp4v file, bf16 build", "file_bytes": os.path.getsize(src)}

    t0 = time.perf_counter()
    frames, fps = vio.read_video_frames(src, -1, -1, -1)
    res["read_video_frames_s"] = time.perf_counter() - t0
    res["decode_frames_per_s"] = frames.shape[0] / res["read_video_frames_s"]
    res["frames_pinned"] = bool(torch.from_numpy(frames).is_pinned())
    pageable = frames.copy()

    model.infer_video_depth(frames[:40], fps, device=dev)  # warm-up
    for name, arr in (("pinned", frames), ("pageable", pageable), ("pinned2", frames)):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        depths, _ = model.infer_video_depth(arr, fps, device=dev)
        res[f"infer_video_depth_{name}_s"] = time.perf_counter() - t0
    res["infer_frames_per_s_pinned"] = a.frames / min(res["infer_video_depth_pinned_s"], res["infer_video_depth_pinned2_s"])
    res["infer_frames_per_s_pageable"] = a.frames / res["infer_video_depth_pageable_s"]

    inferno = np.linspace(0, 1, 256)[:, None].repeat(3, 1) ** np.array([0.5, 1.0, 2.0])  # any table: the work is the same
    for name, kw in (("grey", dict(grayscale=True)), ("colour", dict(colormap=inferno))):
        vio.depth_pictures(depths[:8], device=dev, **kw)
        t0 = time.perf_counter()
        pics = vio.depth_pictures(depths, device=dev, **kw)
        res[f"depth_pictures_{name}_s"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    ref = ((depths - depths.min()) / (depths.max() - depths.min()) * 255).astype(np.uint8)
    ref = (inferno * 255).astype(np.uint8)[ref]
    res["numpy_reference_colour_s"] = time.perf_counter() - t0
    res["pictures_equal_numpy"] = bool(np.array_equal(ref, pics))

    # the two kernels alone, device-resident stack
    lib = L.load()
    d = torch.from_numpy(depths).to(dev)
    n = d.numel()
    partial = torch.empty(int(lib.vda_depth_minmax_workspace_floats()), dtype=torch.float32, device=dev)
    mm = torch.empty(2, dtype=torch.float32, device=dev)
    lut = torch.from_numpy((inferno * 255).astype(np.uint8)).to(dev)
    out = torch.empty(3 * n, dtype=torch.uint8, device=dev)
    s = L.stream_ptr()
    calls = {"minmax": (lambda: lib.vda_depth_minmax(d.data_ptr(), n, partial.data_ptr(), mm.data_ptr(), s), 4 * n),
             "vis_grey": (lambda: lib.vda_depth_vis(d.data_ptr(), n, mm.data_ptr(), None, out.data_ptr(), s), 5 * n),
             "vis_colour": (lambda: lib.vda_depth_vis(d.data_ptr(), n, mm.data_ptr(), lut.data_ptr(), out.data_ptr(), s), 7 * n)}
    peak = 6541.0
    try:
        peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    for name, (fn, nbytes) in calls.items():
        for _ in range(3):
            L.check(fn(), name)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            L.check(fn(), name)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 100
        res[f"kernel_{name}"] = {"us": us, "GB/s": nbytes / us * 1e-3, "frac_of_hbm_peak": nbytes / us * 1e-3 / peak,
                                 "bytes": nbytes}

    t0 = time.perf_counter()
    vio._encode(pics, os.path.join(tmp, "vis.mp4"), fps)
    res["encode_s"] = time.perf_counter() - t0
    res["encode_frames_per_s"] = a.frames / res["encode_s"]
    print(json.dumps(res))


if __name__ == "__main__":
    main()

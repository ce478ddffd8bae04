#!/usr/bin/env python
"""`infere_single_image` (streaming mode) throughput on a synthetic video, one GPU.
Reports frames/s end to end (host uint8 frames in, host fp32 depth out), the steady-state ms per new frame and the
device time of each stage (CUDA events around back-to-back replays of the stage's launch list)."""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--encoder", default="vitl")
    ap.add_argument("--frames", type=int, default=160)
    ap.add_argument("--h", type=int, default=518)
    ap.add_argument("--w", type=int, default=518)
    ap.add_argument("--input-size", type=int, default=518)
    ap.add_argument("--keyframes", default="2,12")
    ap.add_argument("--no-align", action="store_true")
    ap.add_argument("--T", type=int, default=32)
    a = ap.parse_args()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    from bench import build_model
    from scripts.driver_bench import synth
    from video_depth_anything_b200.driver import effective_input_size, network_size
    model = build_model(a.encoder, dev)
    kf = [int(v) for v in a.keyframes.split(",")]
    frames = synth(a.frames, a.h, a.w)
    kw = dict(input_size=a.input_size, device=dev, fp32=False, inference_length=a.T, keyframe_list=kf,
              align_each_new_frame=not a.no_align)
    model.infere_single_image(frames[:a.T + max(kf) + 4], 24, **kw)  # warm-up: weights packed, plans built, graphs captured
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out, _ = model.infere_single_image(frames, 24, **kw)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    # a shorter run isolates the warm-in (T-1 encoder-only steps + the first, large prediction) from the steady state
    n_short = a.T + max(kf) + 4
    t0 = time.perf_counter()
    model.infere_single_image(frames[:n_short], 24, **kw)
    torch.cuda.synchronize()
    dt_short = time.perf_counter() - t0
    steady_ms = (dt - dt_short) / (a.frames - n_short) * 1e3
    nh, nw = network_size(a.h, a.w, effective_input_size(a.h, a.w, a.input_size))
    npreds = (a.T, len(kf) + 2) if not a.no_align else (1,)
    plan = model.stream_plan(a.T, nh, nw, dev, True, a.T + max(kf) - 1, npreds)
    stages = {}
    for name in plan.stages:
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        for _ in range(3):
            plan.run_stage(name)
        ev[0].record()
        for _ in range(10):
            plan.run_stage(name)
        ev[1].record()
        torch.cuda.synchronize()
        stages[name] = {"ms": ev[0].elapsed_time(ev[1]) / 10, "launches": len(plan.stages[name]["ops"])}
    print(json.dumps({"config": f"{a.encoder} infere_single_image, {a.frames} synthetic {a.h}x{a.w} frames, inference_length "
                                f"{a.T}, keyframe_list {kf}, align_each_new_frame {not a.no_align}, bf16 build",
                      "network_input": [nh, nw], "seconds": dt, "frames_per_s_e2e": a.frames / dt,
                      "steady_state_ms_per_frame": steady_ms, "steady_state_frames_per_s": 1e3 / steady_ms,
                      "out_shape": list(out.shape), "h2d_bytes": int(frames.nbytes), "d2h_bytes": int(out.nbytes),
                      "stage_device_ms": stages}))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""`infer_video_depth` throughput on synthetic videos (BASELINE.json configs 3 and 4), one GPU or under torchrun.
Reports frames/s end to end (host uint8 frames in, host fp32 depth out) and the per-stage device times."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def synth(n, h, w, seed=0):
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (n // 8 + 2, h // 8 + 1, w // 8 + 1, 3), dtype=np.uint8)
    return np.repeat(np.repeat(np.repeat(base, 8, 0), 8, 1), 8, 2)[:n, :h, :w].copy()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--encoder", default="vits")
    ap.add_argument("--frames", type=int, default=1024)
    ap.add_argument("--h", type=int, default=518)
    ap.add_argument("--w", type=int, default=518)
    ap.add_argument("--input-size", type=int, default=518)
    ap.add_argument("--repeat", type=int, default=2)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    from bench import build_model
    from video_depth_anything_b200.driver import (WindowedDepthDriver, effective_input_size, network_size, window_index_map)
    model = build_model(a.encoder, dev)
    frames = synth(a.frames, a.h, a.w)
    drv = WindowedDepthDriver(model, device=dev, fp32=False, group=dist.group.WORLD if dist is not None else None)
    drv.run(frames[:44 * world], a.input_size)  # warm-up: weights packed, plan built
    torch.cuda.synchronize()
    best = None
    for _ in range(a.repeat):
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        out = drv.run(frames, a.input_size)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    if rank == 0:
        isz = effective_input_size(a.h, a.w, a.input_size)
        nh, nw = network_size(a.h, a.w, isz)
        nwin = window_index_map(a.frames).shape[0]
        print(json.dumps({"config": f"{a.encoder} infer_video_depth, {a.frames} synthetic {a.h}x{a.w} frames, input_size {a.input_size}",
                          "n_gpus": world, "network_input": [nh, nw], "windows": nwin, "seconds": best,
                          "frames_per_s_e2e": a.frames / best, "ms_per_window": best / nwin * 1e3,
                          "out_shape": list(out.shape) if out is not None else None,
                          "h2d_bytes": int(frames.nbytes), "d2h_bytes": int(out.nbytes) if out is not None else 0,
                          "stages_s": getattr(drv, "last_stage_s", None)}))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

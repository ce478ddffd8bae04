#!/usr/bin/env python
"""Run under torchrun (one rank per GPU): `infer_video_depth` with windows sharded over the ranks must return, on rank 0,
exactly what a single rank computes; prints timings of both.  vits, synthetic video."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=200)
    ap.add_argument("--h", type=int, default=270)
    ap.add_argument("--w", type=int, default=480)
    ap.add_argument("--input-size", type=int, default=518)
    ap.add_argument("--encoder", default="vits")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from bench import build_model
    from video_depth_anything_b200.driver import WindowedDepthDriver
    model = build_model(a.encoder, dev)
    rng = np.random.default_rng(0)
    base = rng.integers(0, 256, (a.frames // 8 + 2, a.h // 8 + 1, a.w // 8 + 1, 3), dtype=np.uint8)
    frames = np.repeat(np.repeat(np.repeat(base, 8, 0), 8, 1), 8, 2)[:a.frames, :a.h, :a.w].copy()  # blocky, non-degenerate
    drv = WindowedDepthDriver(model, device=dev, fp32=False, group=dist.group.WORLD)
    drv.run(frames[:22 * 2 * world], a.input_size)  # warm-up (plans, weights, NCCL)
    drv._shared_result(frames.shape[0] * a.h * a.w * 4)
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    out = drv.run(frames, a.input_size)
    torch.cuda.synchronize()
    dist.barrier()
    t_sharded = time.perf_counter() - t0
    stages = [None] * world
    dist.all_gather_object(stages, drv.last_stage_s)
    if rank == 0:
        out = np.array(out, copy=True)
        solo = WindowedDepthDriver(model, device=dev, fp32=False)
        solo.run(frames[:44], a.input_size, fresh=False)
        solo._pinned_staging(frames.shape[0] * a.h * a.w * 4)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ref = solo.run(frames, a.input_size, fresh=False)
        torch.cuda.synchronize()
        t_solo = time.perf_counter() - t0
        same = bool(np.array_equal(out, ref))
        print(json.dumps({"world": world, "frames": a.frames, "frame_hw": [a.h, a.w], "encoder": a.encoder,
                          "sharded_s": t_sharded, "single_rank_s": t_solo, "speedup": t_solo / t_sharded,
                          "frames_per_s_sharded": a.frames / t_sharded, "bit_identical_to_single_rank": same,
                          "max_abs_diff": float(np.abs(out - ref).max()), "stages_per_rank": stages,
                          "single_rank_stages": solo.last_stage_s}))
        assert same, "sharded result differs from the single-rank result"
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

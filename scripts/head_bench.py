#!/usr/bin/env python
"""BASELINE.json config 5: the isolated `DPTHeadTemporal.forward` and its motion modules / temporal attention at
16 / 32 / 64 frames, vits and vitl widths, 37x37 patch grid (518x518 input), bf16 build.  Device times with CUDA events
over back-to-back launches of the launch list (L2 is far smaller than the working set for vitl; stated per line)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def time_ops(plan, ops, iters=5):
    saved = plan.ops
    plan.ops = ops
    for _ in range(2):
        plan._launch_ops()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        plan._launch_ops()
    e1.record()
    torch.cuda.synchronize()
    plan.ops = saved
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--encoders", default="vits,vitl")
    ap.add_argument("--frames", default="16,32,64")
    ap.add_argument("--size", type=int, default=518)
    a = ap.parse_args()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    from bench import CFG
    from video_depth_anything_b200 import lib as L
    from video_depth_anything_b200.video_depth import VideoDepthAnything
    lib = L.load()
    for enc in a.encoders.split(","):
        for T in [int(t) for t in a.frames.split(",")]:
            torch.manual_seed(0)
            m = VideoDepthAnything(**CFG[enc], num_frames=max(T, 32))
            with torch.no_grad():
                for mm in m.head.motion_modules:
                    mm.temporal_transformer.proj_out.weight.normal_(0, 0.02)
            m = m.to(dev).eval()
            plan = m.head_plan(1, T, a.size, a.size, dev, True)
            plan.use_graph = False
            for t in plan.taps:
                t.copy_(torch.randn(t.shape, device=dev).to(t.dtype))
            head_ms = time_ops(plan, plan.ops)
            # the temporal-attention core launches (A18) and whole motion modules (A16-A19: GroupNorm .. proj_out)
            ta = [(fn, args) for fn, args in plan.ops if fn is lib.vda_temporal_attention]
            ta_ms = time_ops(plan, ta)
            ta_bytes = sum(4 * args[3] * args[4] * args[5] * args[6] * 2 for fn, args in ta)  # q,k,v in + o out, bf16
            gn = [i for i, (fn, args) in enumerate(plan.ops) if fn is lib.vda_groupnorm_tokens]
            mm_ms = []
            for k, i0 in enumerate(gn):
                # a module = GroupNorm, proj_in, 2 x (LN, qkv, attention, out), LN, ff1, ff2, proj_out = 14 ops (+2 with rope)
                n_ops = 14 + (2 if plan.pw.rope else 0)
                mm_ms.append(time_ops(plan, plan.ops[i0:i0 + n_ops]))
            print(json.dumps({"config": f"{enc} DPTHeadTemporal.forward alone, {T} frames, {a.size}x{a.size} input "
                                        f"(37x37 patch grid), bf16 build, random taps",
                              "head_ms": head_ms, "head_ms_per_frame": head_ms / T, "launches": len(plan.ops),
                              "motion_module_ms": mm_ms, "temporal_attention_launches": len(ta),
                              "temporal_attention_ms_total": ta_ms,
                              "temporal_attention_GBps": ta_bytes / (ta_ms / 1e3) / 1e9}))
            del plan, m
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

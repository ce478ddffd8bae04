#!/usr/bin/env python
"""Headline benchmark: frames/s of `VideoDepthAnything.forward` on a ViT-L 1x32x518x518 clip, bf16 build, per B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--encoder vitl] [--size 518]

One "step" = one 32-frame window through the whole network (encoder + DPTHeadTemporal).  With N > 1 (launched by
torchrun, one rank per GPU) every rank runs its own window per step -- the window-level data parallelism of
`infer_video_depth` (no collective inside a forward) -- and the time is the max over ranks.  Prints ONE JSON line.

  value      whole-job frames/s with the clip resident in HBM, CUDA-event timed
  e2e        same metric through the public API (`model(x)` under autocast) with pinned host input, H2D and D2H inside
  roofline   tensor-pipe roofline of the dominant kernel (tcgen05 GEMM, encoder qkv/proj/fc1/fc2 launches), timed live
  cpu_baseline  the oracle (CPU restatement of the reference path) on the host cores, bounded sample
--impl reference times that CPU path as its own arm.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = {
    "vits": dict(encoder="vits", features=64, out_channels=[48, 96, 192, 384]),
    "vitl": dict(encoder="vitl", features=256, out_channels=[256, 512, 1024, 1024]),
}
# algorithmic TFLOP per 32-frame clip (counted with hooks on the reference)
CLIP_TFLOP = {("vitl", 518): 44.95, ("vits", 518): 3.881}
T = 32


def measured_traffic():
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture (profiles/traffic.json)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return float(json.load(f)["mean_dram_bytes_per_launch"])
    except (OSError, KeyError, ValueError):
        return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1400.0))), "measured (MEASURED_PEAKS.json, sustained)"
    return 1400.0, "fallback (B200_PROFILING.md, sustained)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx.append(float(c[1]))
            except ValueError:
                continue
            for n, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        if sm:
            sm.sort()
            out = {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def build_model(encoder, device):
    from video_depth_anything_b200.video_depth import VideoDepthAnything
    torch.manual_seed(0)
    m = VideoDepthAnything(**CFG[encoder])
    with torch.no_grad():  # the default init zeroes every motion module's proj_out: make the temporal path live
        for mm in m.head.motion_modules:
            mm.temporal_transformer.proj_out.weight.normal_(0, 0.02)
            mm.temporal_transformer.proj_out.bias.normal_(0, 0.02)
        for blk in m.pretrained.blocks:
            blk.ls1.gamma.uniform_(0.5, 1.5)
            blk.ls2.gamma.uniform_(0.5, 1.5)
    return m.to(device).eval()


def cpu_reference_run(encoder, size, frames, steps, warmup):
    """The reference's own CPU path (oracle port) on all host cores.  Returns (frames/s, seconds per step)."""
    from oracle import model as omodel
    from oracle.weights import make_state_dict
    torch.set_num_threads(os.cpu_count())
    sd = make_state_dict(encoder)
    x = torch.randn(1, frames, 3, size, size, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        for _ in range(warmup):
            omodel.forward(sd, x, encoder)
        t0 = time.perf_counter()
        for _ in range(steps):
            omodel.forward(sd, x, encoder)
        dt = (time.perf_counter() - t0) / steps
    return frames / dt, dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--encoder", default="vitl", choices=["vitl", "vits"])
    ap.add_argument("--size", type=int, default=518)
    ap.add_argument("--cpu-sample-frames", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    enc, size = a.encoder, a.size
    workload = f"Video-Depth-Anything-V2 {enc} forward, synthetic 1x{T}x3x{size}x{size} clip, bf16 autocast build"
    base = {"metric": f"frames/sec {enc} {T}x{size}x{size} bf16 forward (ms_per_step = ms per 32-frame clip)", "unit": "frames/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "data": "synthetic N(0,1) clip, random-init weights (no checkpoints offline)"}

    if a.impl == "reference":
        # the reference is pure PyTorch and cannot travel to the GPU box; its CPU path is timed through the oracle port
        if rank != 0:
            return
        w = max(1, min(a.warmup, 1))
        fps, dt = cpu_reference_run(enc, size, a.cpu_sample_frames, a.steps, w)
        line = dict(base, impl="reference", value=fps, ms_per_step=dt * 1e3, dtype="f32", warmup=w,
                    config={"workload": workload.replace("bf16 autocast build", "fp32 on host CPU"),
                            "sample": f"{a.cpu_sample_frames} of {T} frames per step"},
                    cpu_baseline={"value": fps, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
                                  "sample": f"oracle forward on a 1x{a.cpu_sample_frames}x3x{size}x{size} clip per step, "
                                            f"{a.steps} steps, fp32, torch CPU, {os.cpu_count()} threads"},
                    e2e={"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, gpu_launches=0)
        print(json.dumps(line))
        return

    from video_depth_anything_b200 import lib as L
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    L.require_device(device)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner on stdout when NCCL_DEBUG is set in the environment: keep stdout to the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=device)
    model = build_model(enc, device)
    g = torch.Generator().manual_seed(100 + rank)
    x_host = torch.randn(1, T, 3, size, size, generator=g).pin_memory()
    x_dev = x_host.to(device)
    plan = model.plan(1, T, size, size, device, True)
    plan.x_in.copy_(x_dev.flatten(0, 1))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        plan.launch()
    barrier()

    # ---- timed region: K launches of the whole plan, per-GEMM events on the encoder's tcgen05 GEMMs
    gemm_fn = plan._lib.vda_gemm
    D = plan.pw.D
    enc_gemm = [i for i, (fn, args) in enumerate(plan.ops) if fn is gemm_fn and args[0].use_tc and args[0].taps == 1 and
                args[0].M == plan.BT * (plan.ph * plan.pw_ + 1) and args[0].K in (D, 4 * D)]
    flops_per_step = sum(2.0 * plan.ops[i][1][0].M * plan.ops[i][1][0].N * plan.ops[i][1][0].K for i in enc_gemm)
    evs = []
    sampler = ClockSampler(local_rank) if rank == 0 else None
    st = L.stream_ptr()
    import ctypes
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    enc_set = set(enc_gemm)
    e0.record()
    for _ in range(a.steps):
        for i, (fn, args) in enumerate(plan.ops):
            if i in enc_set:
                s_ev, t_ev = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s_ev.record()
                rc = fn(ctypes.byref(args[0]), st)
                t_ev.record()
                evs.append((s_ev, t_ev))
            elif fn is gemm_fn:
                rc = fn(ctypes.byref(args[0]), st)
            else:
                rc = fn(*args, st)
            if rc != 0:
                raise RuntimeError(plan._lib.vda_last_error().decode())
    e1.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    ms_per_step = ms_total / a.steps
    value = world * T * a.steps / (ms_total / 1e3)
    gemm_ms = sum(s.elapsed_time(t) for s, t in evs)

    # ---- end to end through the public API: pinned host clip -> H2D -> model(x) -> D2H of the depth maps, EVERY step.
    # The three legs are pipelined the way a serving loop runs them (and the way WindowedDepthDriver streams windows):
    # step i+1's upload and step i-1's download ride on side streams under step i's forward.
    up, down = torch.cuda.Stream(device), torch.cuda.Stream(device)
    cur = torch.cuda.current_stream(device)
    xs = [torch.empty_like(x_dev) for _ in range(2)]
    outs = [torch.empty((1, T, size, size), dtype=torch.float32).pin_memory() for _ in range(2)]
    out_host = outs[0]
    up_done = [torch.cuda.Event() for _ in range(2)]
    x_free = [torch.cuda.Event() for _ in range(2)]

    def e2e_steps(n):
        for i in range(n):
            b = i & 1
            with torch.cuda.stream(up):
                up.wait_event(x_free[b])  # the forward that read xs[b] two steps ago is done with it
                xs[b].copy_(x_host, non_blocking=True)
                up_done[b].record(up)
            cur.wait_event(up_done[b])
            d = model(xs[b])
            x_free[b].record(cur)
            with torch.cuda.stream(down):
                down.wait_event(x_free[b])
                outs[b].copy_(d, non_blocking=True)
                d.record_stream(down)
        torch.cuda.synchronize()

    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        e2e_steps(2)
    barrier()
    t0 = time.perf_counter()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        e2e_steps(a.steps)
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=device)
    if dist is not None:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_fps = world * T * a.steps / float(e2e_s.item())

    if rank == 0:
        peak, which = measured_peaks()
        achieved = flops_per_step * a.steps / (gemm_ms / 1e3) / 1e12 if gemm_ms > 0 else 0.0
        clip_tf = CLIP_TFLOP.get((enc, size))
        line = dict(base, value=value, ms_per_step=ms_per_step, dtype="bf16",
                    config={"workload": workload, "frames_per_step_per_gpu": T, "parallelism": f"window-dp{world}",
                            "l2": "per-step working set (~20 GB of activations) is far larger than the 126 MB L2",
                            "algorithmic_tflop_per_clip": clip_tf,
                            "whole_step_tflops": (clip_tf / (ms_per_step / 1e3)) if clip_tf else None},
                    clocks=clocks,
                    e2e={"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": x_host.numel() * 4,
                         "d2h_bytes_per_step": out_host.numel() * 4},
                    gpu_launches=plan.n_launches * a.steps,
                    roofline={"bound": "tensor", "kernel": "gemm_tc2_kernel (tcgen05 cta_group::2 pair GEMM; encoder qkv/proj/fc1/fc2)",
                              "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": measured_traffic(),
                              "peak_source": which, "launches_per_step": len(enc_gemm), "tflop_per_step": flops_per_step / 1e12,
                              "share_of_step": gemm_ms / ms_total})
        if world == 1 and not a.no_cpu_baseline:
            fps, dt = cpu_reference_run(enc, size, a.cpu_sample_frames, 1, 0)
            line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"oracle (CPU restatement of the reference forward) on a 1x{a.cpu_sample_frames}"
                                              f"x3x{size}x{size} clip, fp32, {os.cpu_count()} threads, {dt:.1f} s"}
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

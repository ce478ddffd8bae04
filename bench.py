#!/usr/bin/env python
"""Headline benchmark of the Video-Depth-Anything hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (N > 1)

N = 1   BASELINE.json configs[1]: `VideoDepthAnything.forward` on a ViT-L 1x32x518x518 clip, bf16 build.  One step = one
        32-frame clip through the whole network.  `value` = frames/s with the clip resident in HBM (CUDA events), `e2e` = the
        same through the public API with pinned host input and host output, H2D + D2H inside the timed region, `roofline` =
        the dominant kernel (tcgen05 pair GEMM of the encoder) timed live.  The line also carries `video`: the long-video
        configurations (configs[3]: vitl, 2048 x 720x1280; configs[2]: vits, 1024 x 518x518) through `infer_video_depth`
        on this one GPU -- the single-GPU points of the strong-scaling series -- plus `gpu_eager_baseline` (the reference's
        own torch-eager ops under bf16 autocast on this GPU, via the oracle restatement) and `cpu_baseline`.
N > 1   BASELINE.json configs[3]: ONE vitl video of 2048 720x1280 frames, its 94 windows sharded over the N ranks
        (`infer_video_depth_sharded`: window k on rank k % N, one 11-frame all-gather per round over NCCL, stitching and
        the device->host copy on every rank) -- strong scaling.  One step = one pass over the whole video.  `value` = output
        frames/s with the uint8 frames resident in HBM and the result left in HBM; `e2e` = host frames in, host depth out
        (rank 0 returns the whole [2048, 720, 1280] stack).  `single_gpu_same_workload` is rank 0 alone on the same video in
        the same run, `config3` the vits video, `clip_forward` the N = 1 measurement repeated on every rank (replicas).
--impl reference times the reference's CPU path (oracle port; the reference is pure PyTorch and has no native code to
compile) on the host cores on the same configuration, each step a bounded sample.
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = {
    "vits": dict(encoder="vits", features=64, out_channels=[48, 96, 192, 384]),
    "vitl": dict(encoder="vitl", features=256, out_channels=[256, 512, 1024, 1024]),
}
# algorithmic TFLOP per 32-frame window (counted with hooks on the reference)
CLIP_TFLOP = {("vitl", 518, 518): 44.95, ("vits", 518, 518): 3.881, ("vitl", 518, 924): 88.37, ("vits", 518, 924): 8.462}
T = 32
VIDEO4 = dict(encoder="vitl", frames=2048, fh=720, fw=1280, name="configs[3]")
VIDEO3 = dict(encoder="vits", frames=1024, fh=518, fw=518, name="configs[2]")


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
        sm, mx, pw, reasons = [], [], [], set()
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
            try:
                pw.append(float(c[2]))
            except ValueError:
                pass
            for n, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        if sm:
            sm.sort()
            out = {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
            if pw:   # board power under load: with sw_power_cap active the step time is the step's energy over the cap
                pw.sort()
                out["power_w"] = pw[len(pw) // 2]
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


def make_config(n_gpus, enc="vitl", size=518):
    """The static description of the workload (shared by both arms, so that the driver can compare them)."""
    if n_gpus == 1:
        return {"workload": f"BASELINE.json configs[1]: Video-Depth-Anything-V2 {enc} forward, synthetic 1x{T}x3x{size}x{size} clip, "
                            "bf16 autocast build", "frames_per_step_per_gpu": T, "parallelism": "single GPU",
                "l2": "per-step working set (~20 GB of activations) is far larger than the 126 MB L2",
                "algorithmic_tflop_per_clip": CLIP_TFLOP.get((enc, size, size))}
    v = VIDEO4
    n_win = (v["frames"] + 21) // 22
    return {"workload": f"BASELINE.json {v['name']}: {v['encoder']} infer_video_depth on ONE synthetic {v['frames']}-frame "
                        f"{v['fw']}x{v['fh']} video (input_size 518 -> network 518x924), bf16 build, {n_win} windows of 32 frames "
                        f"sharded over {n_gpus} GPUs (window k on rank k % {n_gpus})",
            "frames_per_step": v["frames"], "windows_per_step": n_win, "parallelism": f"window-sharded over {n_gpus} ranks, "
            "one 11-frame all-gather per round (NCCL), per-rank stitching and device->host copy",
            "l2": "per-window working set (~36 GB of activations) is far larger than the 126 MB L2",
            "algorithmic_tflop_per_window": CLIP_TFLOP[(v["encoder"], 518, 924)]}


def metric_name(n_gpus, enc="vitl", size=518):
    if n_gpus == 1:
        return f"frames/sec {enc} {T}x{size}x{size} bf16 forward (ms_per_step = ms per 32-frame clip)"
    v = VIDEO4
    return (f"frames/sec {v['encoder']} infer_video_depth, {v['frames']} x {v['fw']}x{v['fh']} video, bf16, window-sharded "
            "(output frames per second; ms_per_step = ms per pass over the video)")


# ----------------------------------------------------------------------------------------------- CPU / eager baselines
def cpu_reference_run(encoder, h, w, frames, steps, warmup):
    """The reference's own CPU path (oracle port) on all host cores.  Returns (network frames/s, seconds per step)."""
    from oracle import model as omodel
    from oracle.weights import make_state_dict
    torch.set_num_threads(os.cpu_count())
    sd = make_state_dict(encoder)
    x = torch.randn(1, frames, 3, h, w, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        for _ in range(warmup):
            omodel.forward(sd, x, encoder)
        t0 = time.perf_counter()
        for _ in range(steps):
            omodel.forward(sd, x, encoder)
        dt = (time.perf_counter() - t0) / steps
    return frames / dt, dt


def gpu_eager_run(encoder, size, device, steps=5, warmup=3):
    """The reference's own ops -- torch eager (cuBLASLt / cuDNN / ATen) under bf16 autocast -- on this GPU, through the
    oracle restatement of the reference forward (pinned to the real reference by scripts/make_goldens.py): "the kernel to
    beat on the same box" (BASELINE.md section 4).  Baseline leg only; never part of the product's timed region."""
    from oracle import model as omodel
    from oracle.weights import make_state_dict
    sd = {k: v.to(device) for k, v in make_state_dict(encoder).items()}
    x = torch.randn(1, T, 3, size, size, generator=torch.Generator().manual_seed(0)).to(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        for _ in range(warmup):
            omodel.forward(sd, x, encoder)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            omodel.forward(sd, x, encoder)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    del sd, x
    torch.cuda.empty_cache()
    return T / (ms / 1e3), ms


# ----------------------------------------------------------------------------------------------- synthetic videos
def synthetic_block(n, fh, fw, seed):
    """n distinct uint8 frames [n, fh, fw, 3]: low-pass-filtered noise drifting over time (smooth, non-degenerate depth)."""
    import cv2
    rng = np.random.default_rng(seed)
    gh, gw = fh // 40 + 2, fw // 40 + 2
    a, b = rng.random((gh, gw, 3), dtype=np.float32), rng.random((gh, gw, 3), dtype=np.float32)
    out = np.empty((n, fh, fw, 3), np.uint8)
    for i in range(n):
        t = 0.5 - 0.5 * np.cos(2 * np.pi * i / n)    # periodic: the tiled video has no seam
        img = cv2.resize(a * (1 - t) + b * t, (fw, fh), interpolation=cv2.INTER_CUBIC)
        out[i] = np.clip(img * 255.0, 0, 255).astype(np.uint8)
    return out


def host_video(spec, group, rank):
    """The whole synthetic video as ONE contiguous, page-locked uint8 array [N, H, W, 3]; with a process group it lives in
    shared memory filled by rank 0 (what a multi-process loader would hand over), mapped and page-locked by every rank."""
    n, fh, fw = spec["frames"], spec["fh"], spec["fw"]
    nbytes = n * fh * fw * 3
    if group is None:
        buf = torch.empty((nbytes,), dtype=torch.uint8, pin_memory=True)
        arr = buf.numpy().reshape(n, fh, fw, 3)
        keep = buf
    else:
        import torch.distributed as dist
        from video_depth_anything_b200 import lib as L
        from video_depth_anything_b200.driver import SharedHostBuffer
        keep = SharedHostBuffer(nbytes, group, L.load())
        arr = np.frombuffer(keep.mm, dtype=np.uint8).reshape(n, fh, fw, 3)
    if rank == 0:
        block = synthetic_block(64, fh, fw, 7)
        for lo in range(0, n, 64):
            arr[lo:lo + 64] = block[:min(64, n - lo)]
    if group is not None:
        dist.barrier(group)
    return arr, keep


def run_video(model, spec, device, group, steps, warmup, mode, frames_host=None, frames_dev=None):
    """`steps` timed passes of infer_video_depth over the video.  mode "device": uint8 frames resident in HBM, result left in
    HBM; mode "e2e": host frames in, host depth out.  Returns a dict (rank-local wall seconds are max-reduced by the caller)."""
    import torch.distributed as dist
    from video_depth_anything_b200.driver import WindowedDepthDriver
    drv = WindowedDepthDriver(model, device=device, fp32=False, group=group)
    frames = frames_dev if mode == "device" else frames_host
    short = frames[:3 * 22 * (dist.get_world_size(group) if group is not None else 1)]   # plans, weights, buffers
    drv.run(short, 518, return_device=(mode == "device"), fresh=False)
    if mode == "e2e":   # the cached page-locked result buffer at its full size (page-locking gigabytes takes seconds, once)
        nbytes = spec["frames"] * spec["fh"] * spec["fw"] * 4
        drv._shared_result(nbytes) if group is not None else drv._pinned_staging(nbytes)
    for _ in range(warmup):
        drv.run(frames, 518, return_device=(mode == "device"), fresh=False)
    torch.cuda.synchronize()
    if group is not None:
        dist.barrier(group)
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    out = None
    for _ in range(steps):
        out = drv.run(frames, 518, return_device=(mode == "device"), fresh=False)
    e1.record()
    torch.cuda.synchronize()
    if group is not None:
        dist.barrier(group)
        torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    return {"wall_s": wall, "device_s": e0.elapsed_time(e1) / 1e3, "out": out, "stages": drv.last_stage_s, "driver": drv}


def reduce_max(x, device, group):
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    if group is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def video_numbers(model, spec, device, group, steps, warmup, e2e_steps, frames_host, with_device=True, e2e_warmup=0):
    """Device-resident and end-to-end throughput of one video configuration on the ranks of `group` (None = this GPU)."""
    n = spec["frames"]
    res = {}
    if with_device:
        frames_dev = torch.from_numpy(frames_host).to(device)
        r = run_video(model, spec, device, group, steps, warmup, "device", frames_dev=frames_dev)
        s = max(reduce_max(r["device_s"], device, group), 1e-9) if group is None else reduce_max(r["wall_s"], device, group)
        res["device"] = {"value": n * steps / s, "s_per_pass": s / steps, "steps": steps}
        del frames_dev, r
        torch.cuda.empty_cache()
    r = run_video(model, spec, device, group, e2e_steps, e2e_warmup, "e2e", frames_host=frames_host)
    s = reduce_max(r["wall_s"], device, group)
    res["e2e"] = {"value": n * e2e_steps / s, "s_per_pass": s / e2e_steps, "steps": e2e_steps,
                  "h2d_bytes_per_pass": int(n * spec["fh"] * spec["fw"] * 3), "d2h_bytes_per_pass": int(n * spec["fh"] * spec["fw"] * 4)}
    res["stages_last_pass"] = r["stages"]
    res["_out"] = r["out"]
    return res


# ----------------------------------------------------------------------------------------------- the clip forward (N = 1 headline)
def clip_forward_bench(model, enc, size, device, steps, warmup, rank, local_rank, group):
    import ctypes
    import torch.distributed as dist
    from video_depth_anything_b200 import lib as L
    world = dist.get_world_size(group) if group is not None else 1
    g = torch.Generator().manual_seed(100 + rank)
    x_host = torch.randn(1, T, 3, size, size, generator=g).pin_memory()
    x_dev = x_host.to(device)
    plan = model.plan(1, T, size, size, device, True)
    plan.x_in.copy_(x_dev.flatten(0, 1))

    def barrier():
        if group is not None:
            dist.barrier(group)
        torch.cuda.synchronize()

    for _ in range(warmup):
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
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    enc_set = set(enc_gemm)
    e0.record()
    for _ in range(steps):
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
    ms_total = reduce_max(e0.elapsed_time(e1), device, group)
    gemm_ms = sum(s.elapsed_time(t) for s, t in evs)

    # ---- end to end through the public API: pinned host clip -> H2D -> model(x) -> D2H of the depth maps, EVERY step.
    # The three legs are pipelined the way a serving loop runs them: step i+1's upload and step i-1's download ride on
    # side streams under step i's forward.
    up, down = torch.cuda.Stream(device), torch.cuda.Stream(device)
    cur = torch.cuda.current_stream(device)
    xs = [torch.empty_like(x_dev) for _ in range(2)]
    outs = [torch.empty((1, T, size, size), dtype=torch.float32).pin_memory() for _ in range(2)]
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
        e2e_steps(steps)
    barrier()
    e2e_s = reduce_max(time.perf_counter() - t0, device, group)
    peak, which = measured_peaks()
    achieved = flops_per_step * steps / (gemm_ms / 1e3) / 1e12 if gemm_ms > 0 else 0.0
    clip_tf = CLIP_TFLOP.get((enc, size, size))
    ms_per_step = ms_total / steps
    return {"value": world * T * steps / (ms_total / 1e3), "ms_per_step": ms_per_step, "clocks": clocks,
            "e2e": {"value": world * T * steps / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": x_host.numel() * 4,
                    "d2h_bytes_per_step": outs[0].numel() * 4},
            "gpu_launches": plan.n_launches * steps, "launches_per_window": plan.n_launches,
            "whole_step_tflops": (clip_tf / (ms_per_step / 1e3)) if clip_tf else None,
            "roofline": {"bound": "tensor", "kernel": "gemm_tc2_kernel (tcgen05 cta_group::2 pair GEMM; encoder qkv/proj/fc1/fc2)",
                         "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": measured_traffic(),
                         "peak_source": which, "launches_per_step": len(enc_gemm), "tflop_per_step": flops_per_step / 1e12,
                         "share_of_step": gemm_ms / ms_total}}


_JSON_FD = None


def claim_stdout():
    """The driver reads ONE JSON line from stdout.  Libraries write there too (NCCL prints its version banner on stdout whenever
    NCCL_DEBUG is set, whatever NCCL_DEBUG_FILE says), so the process's fd 1 is pointed at stderr for everybody else and the
    JSON line goes to the saved descriptor."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--encoder", default="vitl", choices=["vitl", "vits"])
    ap.add_argument("--size", type=int, default=518)
    ap.add_argument("--cpu-sample-frames", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-video", action="store_true", help="N = 1: skip the long-video sub-benchmarks")
    ap.add_argument("--no-eager", action="store_true", help="N = 1: skip the torch-eager GPU baseline")
    ap.add_argument("--video-frames", type=int, default=0, help="override the length of the configs[3] video (0 = 2048)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    enc, size = a.encoder, a.size
    if a.video_frames:
        VIDEO4["frames"] = a.video_frames
    n_gpus = a.gpus
    base = {"metric": metric_name(n_gpus, enc, size), "unit": "frames/s", "n_gpus": n_gpus, "steps": a.steps, "warmup": a.warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "data": "synthetic (N(0,1) clip; low-pass-noise uint8 video, 64 distinct frames tiled), random-init weights "
                    "(no checkpoints offline)"}
    config = make_config(n_gpus, enc, size)

    if a.impl == "reference":
        # the reference is pure PyTorch and cannot travel to the GPU box; its CPU path is timed through the oracle port
        if rank != 0:
            return
        w = max(1, min(a.warmup, 1))
        cores = os.cpu_count()
        if n_gpus == 1:
            fps, dt = cpu_reference_run(enc, size, size, a.cpu_sample_frames, a.steps, w)
            sample = (f"each step: oracle forward (CPU restatement of the reference) on {a.cpu_sample_frames} of the clip's {T} frames "
                      f"(1x{a.cpu_sample_frames}x3x{size}x{size}), fp32, torch CPU, {cores} threads; value = frames / time")
            extra = {}
            full_fps, full_dt = cpu_reference_run(enc, size, size, T, 1, 0)   # the whole 32-frame clip, once
            extra["full_clip_once"] = {"frames": T, "seconds": full_dt, "frames_per_s": full_fps}
        else:
            v = VIDEO4
            nf = 4
            net_fps, dt = cpu_reference_run(v["encoder"], 518, 924, nf, a.steps, w)
            fps = net_fps * 22.0 / 32.0   # a window runs 32 network frames for 22 new output frames
            sample = (f"each step: oracle forward on {nf} network frames at the video's network size 518x924, fp32, torch CPU, "
                      f"{cores} threads; value = network frames/s x 22/32 (output frames per window); pre-processing and "
                      "stitching (host numpy in the reference) not included")
            extra = {}
        line = dict(base, impl="reference", value=fps, ms_per_step=dt * 1e3, dtype="f32", warmup=w, config=config,
                    cpu_baseline=dict({"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample}, **extra),
                    e2e={"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, gpu_launches=0)
        emit(line)
        return

    from video_depth_anything_b200 import lib as L
    from video_depth_anything_b200.engine import active_toggles
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    L.require_device(device)
    dist = group = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
        group = dist.group.WORLD
    model = build_model(enc, device)
    clip = clip_forward_bench(model, enc, size, device, a.steps if world == 1 else min(a.steps, 10), max(a.warmup, 3), rank,
                              local_rank, group)
    toggles = active_toggles()

    if world == 1:
        line = dict(base, value=clip["value"], ms_per_step=clip["ms_per_step"], dtype="bf16", config=config, clocks=clip["clocks"],
                    e2e=clip["e2e"], gpu_launches=clip["gpu_launches"], roofline=clip["roofline"],
                    derived={"whole_step_tflops": clip["whole_step_tflops"], "launches_per_window": clip["launches_per_window"]},
                    env_toggles=toggles)
        if not a.no_eager:
            fps, ms = gpu_eager_run(enc, size, device)
            line["gpu_eager_baseline"] = {"value": fps, "unit": "frames/s", "ms_per_clip": ms, "ratio_ours_over_eager": clip["value"] / fps,
                                          "what": "the reference's torch-eager ops (oracle restatement of the reference forward: cuBLASLt / "
                                                  "cuDNN / ATen kernels) under torch.autocast('cuda', bfloat16) on the same clip and GPU, "
                                                  "CUDA events, 3 warm-up + 5 timed forwards"}
        if not a.no_video:
            video = {}
            for key, spec in (("config4", VIDEO4), ("config3", VIDEO3)):
                m = model if spec["encoder"] == enc else build_model(spec["encoder"], device)
                model._plans.clear()
                torch.cuda.empty_cache()
                frames_host, keep = host_video(spec, None, 0)
                r = video_numbers(m, spec, device, None, 1, 0, 1, frames_host)
                r.pop("_out")
                video[key] = dict(r, workload=f"{spec['encoder']} infer_video_depth, {spec['frames']} x {spec['fw']}x{spec['fh']}, 1 GPU",
                                  unit="output frames/s")
                del frames_host, keep, m
                torch.cuda.empty_cache()
            line["video"] = video
        if not a.no_cpu_baseline:
            fps, dt = cpu_reference_run(enc, size, size, a.cpu_sample_frames, 1, 0)
            line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"oracle (CPU restatement of the reference forward) on {a.cpu_sample_frames} of the clip's "
                                              f"{T} frames (1x{a.cpu_sample_frames}x3x{size}x{size}), fp32, {os.cpu_count()} threads, {dt:.1f} s"}
        emit(line)
        return

    # ------------------------------------------------------------------ N > 1: the window-sharded long video (configs[3])
    spec = VIDEO4
    model._plans.clear()
    torch.cuda.empty_cache()
    frames_host, keep = host_video(spec, group, rank)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    r = video_numbers(model, spec, device, group, a.steps, a.warmup, min(a.steps, 5), frames_host, e2e_warmup=1)
    clocks = sampler.stop() if sampler else None
    sharded_out = r.pop("_out")
    stages_all = [None] * world
    dist.all_gather_object(stages_all, r.pop("stages_last_pass"), group=group)
    # rank 0 alone on the same video, same run: the single-GPU point of the strong-scaling series, and the bit-identity check
    solo = None
    if rank == 0:
        keep_copy = np.array(sharded_out, copy=True)
        s = video_numbers(model, spec, device, None, 1, 0, 1, frames_host, with_device=False)
        solo = {"e2e": s["e2e"], "bit_identical_to_sharded": bool(np.array_equal(keep_copy, s.pop("_out")))}
        del keep_copy
    dist.barrier(group)
    # configs[2]: vits, 1024 x 518x518
    m3 = build_model(VIDEO3["encoder"], device)
    f3, keep3 = host_video(VIDEO3, group, rank)
    r3 = video_numbers(m3, VIDEO3, device, group, 3, 1, 3, f3, e2e_warmup=1)
    r3.pop("_out")
    r3.pop("stages_last_pass")
    if rank == 0:
        n_win = config["windows_per_step"]
        st0 = stages_all[0] or {}
        launches = (clip["launches_per_window"] + 12 + 3 * world) * n_win * a.steps
        # BASELINE.json's metric is clips at 1 GPU and long videos at 2 / 4 / 8: the N = 1 line times configs[1], this line
        # configs[3], so v_N / (N v_1) across the two lines compares different workloads (a 518x924 window costs 2.1x a
        # 518x518 clip).  The single-GPU figure for THIS video, measured in this run, is single_gpu_same_workload.
        config = dict(config, n1_line_workload="configs[1] (one 32-frame 518x518 clip forward); same-video single-GPU run: "
                                                 "single_gpu_same_workload", replicated_clip_forward="clip_forward (the N = 1 "
                                                 "workload on every rank at once)")
        line = dict(base, value=r["device"]["value"], ms_per_step=r["device"]["s_per_pass"] * 1e3, dtype="bf16", config=config,
                    clocks=clocks,
                    e2e={"value": r["e2e"]["value"], "unit": "frames/s", "h2d_bytes_per_step": r["e2e"]["h2d_bytes_per_pass"],
                         "d2h_bytes_per_step": r["e2e"]["d2h_bytes_per_pass"], "ms_per_step": r["e2e"]["s_per_pass"] * 1e3,
                         "steps": r["e2e"]["steps"]},
                    gpu_launches=launches, roofline=clip["roofline"],
                    stages={"what": "CUDA-event seconds of the last end-to-end pass, summed over the rank's rounds (forward on the main "
                                    "stream; exchange = NCCL all-gather incl. waiting for the slowest rank, stitch and d2h on the side "
                                    "stream, overlapped with the next forward)", "per_rank": stages_all,
                            "nccl_bytes_sent_per_window": 11 * spec["fh"] * spec["fw"] * 4,
                            "nccl_bytes_received_per_rank_per_round": st0.get("exchange_bytes_per_round")},
                    single_gpu_same_workload=solo, config3=dict(r3, workload="BASELINE.json configs[2]: vits infer_video_depth, 1024 x "
                                                                "518x518, sharded the same way", unit="output frames/s"),
                    clip_forward={"value": clip["value"], "ms_per_step": clip["ms_per_step"], "e2e": clip["e2e"],
                                  "what": "the N = 1 workload (vitl 1x32x518x518 forward) on every rank at once: replicas, no exchange"},
                    env_toggles=toggles)
        emit(line)
    dist.barrier(group)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""The sliding-window driver: `VideoDepthAnything.infer_video_depth` (reference video_depth.py:329-417).

What the reference does on the host with cv2 / numpy (pre-processing 32x per window, per-frame D2H, numpy
least squares and blending) runs here on the device, asynchronously on one stream: every source frame is
pre-processed once, windows gather their 32 inputs by index (the window contents are pure index arithmetic,
video_depth.py:350-364), and the sequential scale/shift recurrence keeps (scale, shift) in device memory, so there
is exactly one device->host copy: the final [N, H, W] stack.

Window-level data parallelism: with `torch.distributed` initialised, windows are dealt to
ranks in contiguous blocks; every rank runs its forwards, rank 0 gathers the raw per-window depth maps (NCCL) and
runs the cheap sequential stitching pass.  No collective runs inside a forward.
"""
from __future__ import annotations

import queue
import threading
import time

import numpy as np
import torch

from . import lib as L

INFER_LEN = 32
OVERLAP = 10
KEYFRAMES = [0, 12, 24, 25, 26, 27, 28, 29, 30, 31]
INTERP_LEN = 8
FRAME_STEP = INFER_LEN - OVERLAP
ALIGN_LEN = OVERLAP - INTERP_LEN


def effective_input_size(frame_h: int, frame_w: int, input_size: int) -> int:
    """video_depth.py:330-334."""
    ratio = max(frame_h, frame_w) / min(frame_h, frame_w)
    if ratio > 1.78:
        input_size = int(input_size * 1.777 / ratio)
        input_size = round(input_size / 14) * 14
    return input_size


def network_size(frame_h: int, frame_w: int, input_size: int, multiple: int = 14):
    """`Resize.get_size`, keep_aspect_ratio + 'lower_bound' + ensure_multiple_of=14 (util/transform.py:51-107)."""

    def constrain(x, min_val):
        y = int(np.round(x / multiple) * multiple)
        if y < min_val:
            y = int(np.ceil(x / multiple) * multiple)
        return y

    scale_h, scale_w = input_size / frame_h, input_size / frame_w
    if scale_w > scale_h:
        scale_h = scale_w
    else:
        scale_w = scale_h
    return constrain(scale_h * frame_h, input_size), constrain(scale_w * frame_w, input_size)


def window_index_map(n_frames: int) -> np.ndarray:
    """[n_windows, 32] source-frame index of every window slot (video_depth.py:350-364,374)."""
    append = (FRAME_STEP - (n_frames % FRAME_STEP)) % FRAME_STEP + (INFER_LEN - FRAME_STEP)
    padded = np.concatenate([np.arange(n_frames), np.full(append, n_frames - 1)]).astype(np.int64)
    starts = range(0, n_frames, FRAME_STEP)
    idx = np.stack([padded[s:s + INFER_LEN] for s in starts])
    for k in range(1, idx.shape[0]):
        idx[k, :OVERLAP] = idx[k - 1, KEYFRAMES]
    return idx


def shard_windows(n_windows: int, world: int):
    """Contiguous blocks of windows per rank; [(first, count)] * world."""
    per = (n_windows + world - 1) // world
    out = []
    for r in range(world):
        lo = min(r * per, n_windows)
        out.append((lo, max(0, min(per, n_windows - lo))))
    return out


class WindowedDepthDriver:
    def __init__(self, model, device="cuda", fp32=False, skip_tmp_block=False, upload_chunk=64, distributed=True, streaming=True):
        self.model = model
        self.device = torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        L.require_device(self.device)
        self.fp32 = fp32
        self.skip_tmp_block = skip_tmp_block
        self.upload_chunk = upload_chunk
        self.distributed = distributed  # False: ignore an initialised process group (single-rank reference runs)
        self.streaming = streaming      # single rank: pipeline upload / forward / stitch / download window by window
        self.lib = L.load()
        self.last_scale_shift = None
        self.last_stage_s = None

    # ------------------------------------------------------------------ stages
    def preprocess(self, frames, ids, nh, nw):
        """Source frames `ids` (sorted unique) -> {id: row} and device tensor [len(ids), 3, nh, nw] fp32
        (util/transform.py:109-158 + video_depth.py:361, once per source frame)."""
        fh, fw = frames[0].shape[:2]
        out = torch.empty((len(ids), 3, nh, nw), dtype=torch.float32, device=self.device)
        st = L.stream_ptr()
        for c0 in range(0, len(ids), self.upload_chunk):
            chunk = ids[c0:c0 + self.upload_chunk]
            host = np.ascontiguousarray(np.stack([np.asarray(frames[int(i)]) for i in chunk]))
            if host.dtype != np.uint8:
                raise TypeError("frames must be uint8 [N,H,W,3]")
            dev = torch.from_numpy(host).to(self.device, non_blocking=True)
            L.check(self.lib.vda_preprocess_u8(dev.data_ptr(), out[c0].data_ptr(), len(chunk), fh, fw, nh, nw, st),
                    "vda_preprocess_u8")
            del dev
        return {int(i): r for r, i in enumerate(ids)}, out

    def forward_windows(self, pre, rowmap, idx_rows, nh, nw, fh, fw):
        """Run the network on the given windows; returns raw depth [n_win*32, fh, fw] fp32 on device."""
        bf16 = not self.fp32
        plan = self.model.plan(1, INFER_LEN, nh, nw, self.device, bf16, self.skip_tmp_block)
        raw = torch.empty((len(idx_rows) * INFER_LEN, fh, fw), dtype=torch.float32, device=self.device)
        st = L.stream_ptr()
        sel_all = torch.as_tensor([[rowmap[int(i)] for i in row] for row in idx_rows], dtype=torch.int64).to(self.device)
        for k in range(len(idx_rows)):
            torch.index_select(pre, 0, sel_all[k], out=plan.x_in)
            plan.launch()
            L.check(self.lib.vda_resize_depth(plan.depth.data_ptr(), L.F32, raw[k * INFER_LEN].data_ptr(), INFER_LEN,
                                              nh, nw, fh, fw, st), "vda_resize_depth")
        return raw

    def stitch(self, raw, n_frames):
        """Scale/shift alignment + overlap blend over all windows (video_depth.py:379-417), on device."""
        n_win = raw.shape[0] // INFER_LEN
        fh, fw = raw.shape[1:]
        fe = fh * fw
        total = INFER_LEN + FRAME_STEP * (n_win - 1)
        out = torch.empty((total, fh, fw), dtype=torch.float32, device=self.device)
        out[:INFER_LEN].copy_(raw[:INFER_LEN])
        ref = torch.stack([raw[KEYFRAMES[0]], raw[KEYFRAMES[1]]]).contiguous()   # frame 0 of the video + slot 12
        st_all = torch.zeros((n_win, 2), dtype=torch.float32, device=self.device)
        st_all[0, 0] = 1.0
        scratch = torch.zeros(5, dtype=torch.float64, device=self.device)
        s = L.stream_ptr()
        pos = INFER_LEN
        for k in range(1, n_win):
            cur = raw[k * INFER_LEN:(k + 1) * INFER_LEN]
            stp = st_all[k].data_ptr()
            L.check(self.lib.vda_scale_shift(cur.data_ptr(), ref.data_ptr(), ALIGN_LEN * fe, scratch.data_ptr(), stp, 3, s),
                    "vda_scale_shift")
            L.check(self.lib.vda_blend(out[pos - INTERP_LEN].data_ptr(), cur[ALIGN_LEN].data_ptr(), stp, INTERP_LEN, fe, s),
                    "vda_blend")
            L.check(self.lib.vda_scale_clamp(cur[OVERLAP].data_ptr(), out[pos].data_ptr(), stp, FRAME_STEP * fe, s),
                    "vda_scale_clamp")
            L.check(self.lib.vda_scale_clamp(cur[KEYFRAMES[1]].data_ptr(), ref[1].data_ptr(), stp, fe, s), "vda_scale_clamp")
            pos += FRAME_STEP
        self.last_scale_shift = st_all
        return out[:n_frames]

    # ------------------------------------------------------------------ whole job
    def run(self, frames, input_size=518, return_device=False):
        n = frames.shape[0]
        fh, fw = frames[0].shape[:2]
        isz = effective_input_size(fh, fw, input_size)
        nh, nw = network_size(fh, fw, isz)
        idx = window_index_map(n)
        dist = torch.distributed
        world = dist.get_world_size() if self.distributed and dist.is_available() and dist.is_initialized() else 1
        rank = dist.get_rank() if world > 1 else 0
        first, count = shard_windows(idx.shape[0], world)[rank]
        mine = idx[first:first + count]
        if world == 1 and self.streaming:
            with torch.cuda.device(self.device), torch.no_grad():
                return self._run_streaming(frames, n, idx, nh, nw, fh, fw, return_device)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        with torch.cuda.device(self.device), torch.no_grad():
            ev[0].record()
            if count > 0:
                ids = np.unique(mine)
                rowmap, pre = self.preprocess(frames, ids, nh, nw)
                ev[1].record()
                raw = self.forward_windows(pre, rowmap, mine, nh, nw, fh, fw)
                del pre
            else:
                ev[1].record()
                raw = torch.empty((0, fh, fw), dtype=torch.float32, device=self.device)
            ev[2].record()
            if world > 1:
                raw = self._gather_raw(raw, idx.shape[0], world, rank, fh, fw)
                if rank != 0:
                    return None
            ev[3].record()
            depth = self.stitch(raw, n)
            ev[4].record()
            if return_device:
                return depth
            t0 = time.perf_counter()
            host = self._download(depth)
            # device-side stage times of this call (CUDA events on the launching stream) + the final D2H wall time
            self.last_stage_s = {"upload_preprocess": ev[0].elapsed_time(ev[1]) / 1e3, "forward_windows": ev[1].elapsed_time(ev[2]) / 1e3,
                                 "gather": ev[2].elapsed_time(ev[3]) / 1e3, "stitch": ev[3].elapsed_time(ev[4]) / 1e3,
                                 "d2h": time.perf_counter() - t0}
            return host

    # ------------------------------------------------------------------ single-rank streaming pipeline
    def _upload_preprocess(self, frames, lo, hi, pre, nh, nw):
        """Source frames [lo, hi) -> pre[lo:hi] (pre-processing runs once per source frame)."""
        fh, fw = frames[0].shape[:2]
        st = L.stream_ptr()
        for c0 in range(lo, hi, self.upload_chunk):
            c1 = min(hi, c0 + self.upload_chunk)
            if isinstance(frames, np.ndarray) and frames.flags.c_contiguous:
                host = frames[c0:c1]
            else:
                host = np.ascontiguousarray(np.stack([np.asarray(frames[i]) for i in range(c0, c1)]))
            if host.dtype != np.uint8:
                raise TypeError("frames must be uint8 [N,H,W,3]")
            dev = torch.from_numpy(host).to(self.device, non_blocking=True)
            L.check(self.lib.vda_preprocess_u8(dev.data_ptr(), pre[c0].data_ptr(), c1 - c0, fh, fw, nh, nw, st), "vda_preprocess_u8")
            del dev

    def _run_streaming(self, frames, n, idx, nh, nw, fh, fw, return_device):
        """Window k's upload + pre-processing is issued while window k-1 computes (the host copy overlaps the queued
        kernels), every window is stitched as soon as it is done (the (scale, shift) recurrence is sequential in window
        order anyway, video_depth.py:379-413), and frames that can no longer change are copied to pinned host memory on
        a side stream while later windows run.  Same arithmetic, same order, as the batch path."""
        bf16 = not self.fp32
        plan = self.model.plan(1, INFER_LEN, nh, nw, self.device, bf16, self.skip_tmp_block)
        n_win = idx.shape[0]
        fe = fh * fw
        total = INFER_LEN + FRAME_STEP * (n_win - 1)
        s = L.stream_ptr()
        main = torch.cuda.current_stream()
        pre = torch.empty((n, 3, nh, nw), dtype=torch.float32, device=self.device)
        out = torch.empty((total, fh, fw), dtype=torch.float32, device=self.device)
        raw = torch.empty((2, INFER_LEN, fh, fw), dtype=torch.float32, device=self.device)  # ring: window k -> slot k & 1
        ref = torch.empty((2, fh, fw), dtype=torch.float32, device=self.device)
        st_all = torch.zeros((n_win, 2), dtype=torch.float32, device=self.device)
        st_all[0, 0] = 1.0
        scratch = torch.zeros(5, dtype=torch.float64, device=self.device)
        sel_all = torch.as_tensor(idx, dtype=torch.int64).to(self.device)
        host = None
        copy_stream = None
        if not return_device:
            # device -> pinned staging (cached on the model: page-locking 1 GB costs ~0.5 s) -> the caller's pageable
            # array, the second hop on a worker thread while the GPU keeps computing
            host = self._pinned_staging(n * fe * 4).view(torch.float32)[:n * fe].view(n, fh, fw)
            host_np = host.numpy()
            result = np.empty((n, fh, fw), dtype=np.float32)
            copy_stream = torch.cuda.Stream(device=self.device)
            jobs = queue.Queue()

            def drain():
                while True:
                    job = jobs.get()
                    if job is None:
                        return
                    ev, lo, hi = job
                    ev.synchronize()
                    np.copyto(result[lo:hi], host_np[lo:hi])

            worker = threading.Thread(target=drain, daemon=True)
            worker.start()
        uploaded = 0
        copied = 0
        t_begin = time.perf_counter()
        for k in range(n_win):
            need = int(idx[k].max()) + 1
            if need > uploaded:
                self._upload_preprocess(frames, uploaded, need, pre, nh, nw)
                uploaded = need
            cur = raw[k & 1]
            torch.index_select(pre, 0, sel_all[k], out=plan.x_in)
            plan.launch()
            L.check(self.lib.vda_resize_depth(plan.depth.data_ptr(), L.F32, cur.data_ptr(), INFER_LEN, nh, nw, fh, fw, s),
                    "vda_resize_depth")
            if k == 0:
                out[:INFER_LEN].copy_(cur)
                ref[0].copy_(cur[KEYFRAMES[0]])
                ref[1].copy_(cur[KEYFRAMES[1]])
                pos = INFER_LEN
            else:
                stp = st_all[k].data_ptr()
                L.check(self.lib.vda_scale_shift(cur.data_ptr(), ref.data_ptr(), ALIGN_LEN * fe, scratch.data_ptr(), stp, 3, s),
                        "vda_scale_shift")
                L.check(self.lib.vda_blend(out[pos - INTERP_LEN].data_ptr(), cur[ALIGN_LEN].data_ptr(), stp, INTERP_LEN, fe, s),
                        "vda_blend")
                L.check(self.lib.vda_scale_clamp(cur[OVERLAP].data_ptr(), out[pos].data_ptr(), stp, FRAME_STEP * fe, s),
                        "vda_scale_clamp")
                L.check(self.lib.vda_scale_clamp(cur[KEYFRAMES[1]].data_ptr(), ref[1].data_ptr(), stp, fe, s), "vda_scale_clamp")
                pos += FRAME_STEP
            if host is not None:
                final = n if k == n_win - 1 else min(n, pos - INTERP_LEN)  # the last 8 frames are still to be cross-faded
                if final > copied:
                    done = torch.cuda.Event()
                    done.record(main)
                    copy_stream.wait_event(done)
                    with torch.cuda.stream(copy_stream):
                        host[copied:final].copy_(out[copied:final], non_blocking=True)
                        landed = torch.cuda.Event()
                        landed.record(copy_stream)
                    jobs.put((landed, copied, final))
                    copied = final
        self.last_scale_shift = st_all
        if host is None:
            return out[:n]
        jobs.put(None)
        worker.join()
        main.synchronize()
        self.last_stage_s = {"streaming_total": time.perf_counter() - t_begin}
        return result

    def _download(self, depth, chunk_frames=32, workers=4):
        """Device [n, H, W] fp32 -> a fresh pageable numpy array.  A plain `.cpu()` into pageable memory moves ~2 GB/s (the
        runtime stages through small internal buffers and the page faults of the new array land on one thread); here the
        stack goes to the cached page-locked staging buffer in chunks on a side stream at PCIe rate while worker threads
        copy finished chunks out (numpy releases the GIL for the memcpy, the first-touch faults spread over the threads)."""
        n, fh, fw = depth.shape
        if n == 0:
            return np.empty((0, fh, fw), dtype=np.float32)
        stage = self._pinned_staging(depth.numel() * 4).view(torch.float32)[:depth.numel()].view(n, fh, fw)
        stage_np = stage.numpy()
        result = np.empty((n, fh, fw), dtype=np.float32)
        jobs = queue.Queue()

        def drain():
            while True:
                job = jobs.get()
                if job is None:
                    return
                ev, lo, hi = job
                ev.synchronize()
                np.copyto(result[lo:hi], stage_np[lo:hi])

        pool = [threading.Thread(target=drain, daemon=True) for _ in range(workers)]
        for t in pool:
            t.start()
        main = torch.cuda.current_stream()
        side = torch.cuda.Stream(device=self.device)
        ready = torch.cuda.Event()
        ready.record(main)
        side.wait_event(ready)
        with torch.cuda.stream(side):
            for lo in range(0, n, chunk_frames):
                hi = min(n, lo + chunk_frames)
                stage[lo:hi].copy_(depth[lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(side)
                jobs.put((ev, lo, hi))
        for _ in pool:
            jobs.put(None)
        for t in pool:
            t.join()
        depth.record_stream(side)
        return result

    def _pinned_staging(self, nbytes):
        """Page-locked staging buffer, cached on the model and grown on demand."""
        buf = getattr(self.model, "_vda_pinned_staging", None)
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty((nbytes,), dtype=torch.uint8, pin_memory=True)
            self.model._vda_pinned_staging = buf
        return buf

    def _gather_raw(self, raw, n_win, world, rank, fh, fw):
        return gather_window_blocks(raw, n_win, world, rank)


def gather_window_blocks(raw, n_win, world, rank):
    """Raw depth blocks of every rank -> rank 0, point-to-point (NCCL over NVLink on GPUs; the same code runs over gloo
    in the CPU tests).  Block sizes follow from `shard_windows`, so no size exchange is needed.  Returns the full
    [n_win*32, H, W] stack on rank 0 and the local block elsewhere."""
    dist = torch.distributed
    shards = shard_windows(n_win, world)
    if rank == 0:
        full = torch.empty((n_win * INFER_LEN,) + tuple(raw.shape[1:]), dtype=raw.dtype, device=raw.device)
        full[:raw.shape[0]].copy_(raw)
        reqs = []
        for r in range(1, world):
            lo, cnt = shards[r]
            if cnt:
                reqs.append(dist.irecv(full[lo * INFER_LEN:(lo + cnt) * INFER_LEN], src=r))
        for q in reqs:
            q.wait()
        return full
    if raw.shape[0]:
        dist.send(raw.contiguous(), dst=0)
    return raw

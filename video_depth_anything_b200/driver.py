"""The sliding-window driver: `VideoDepthAnything.infer_video_depth` (reference video_depth.py:329-417).

What the reference does on the host with cv2 / numpy (pre-processing 32x per window, per-frame D2H, numpy least squares
and blending) runs here on the device, as a pipeline over three streams: every source frame is uploaded and
pre-processed once (upload stream), windows gather their 32 inputs by pointer (the window contents are pure index
arithmetic, video_depth.py:350-364) and run the network (main stream), and each finished window is aligned, cross-faded
and copied to the host while the next one computes (stitch stream).

Window-level data parallelism, opt-in through a process group: window k belongs to rank
k % world, so every `world` consecutive windows form a ROUND that the ranks compute at the same time.  No collective runs
inside a forward.  After its forward a rank shares 11 raw frames of its window with the others (one all-gather per round
over NCCL / NVLink): slots 0, 1 (the two frames of the scale/shift fit), slot 12 (the next window's alignment target) and
slots 24..31 (the tail its successor cross-fades with).  Every rank then replays the scalar (scale, shift) recurrence of
the round (video_depth.py:393-395,409-413) with a deterministic reduction -- bit-identical on all ranks and to a
one-rank run -- finalises the 22 frames its own window contributes, and copies them over ITS OWN PCIe link into a host
buffer shared by the ranks of the node.  The result is bit-identical to a single-rank run.
"""
from __future__ import annotations

import mmap
import os
import queue
import threading
import time
import warnings

import numpy as np
import torch

from . import lib as L

INFER_LEN = 32
OVERLAP = 10
KEYFRAMES = [0, 12, 24, 25, 26, 27, 28, 29, 30, 31]
INTERP_LEN = 8
FRAME_STEP = INFER_LEN - OVERLAP
ALIGN_LEN = OVERLAP - INTERP_LEN
# what a window shares with the other ranks: the fit frames (0, 1), the next alignment target (12), the cross-fade tail
XCHG_SLOTS = list(range(ALIGN_LEN)) + [KEYFRAMES[1]] + list(range(INFER_LEN - INTERP_LEN, INFER_LEN))
XCHG_FIT, XCHG_REF, XCHG_TAIL = slice(0, ALIGN_LEN), ALIGN_LEN, slice(ALIGN_LEN + 1, ALIGN_LEN + 1 + INTERP_LEN)


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


def rank_windows(n_windows: int, world: int, rank: int):
    """Windows of `rank`: k = rank, rank + world, ... (round j holds windows j*world .. j*world + world - 1)."""
    return list(range(rank, n_windows, world))


def final_range(k: int, n_windows: int):
    """Output frames [lo, hi) that window k finalises (video_depth.py:379-415): window 0 its first 24 frames, window k >= 1
    the 8 frames it cross-fades with its predecessor's tail (22k+2 ..) and its next 14; the last window also its tail."""
    lo = 0 if k == 0 else FRAME_STEP * k + ALIGN_LEN
    hi = FRAME_STEP * k + INFER_LEN - INTERP_LEN
    if k == n_windows - 1:
        hi += INTERP_LEN
    return lo, hi


POOL_GROUPS = 3   # the frame pool is a ring of three window-sized groups behind one permanent row for frame 0


def upload_runs(idx_rows):
    """For a rank's windows in processing order: the source frames each one has to upload, as runs of consecutive ids, and
    the pool row of each of the window's 32 slots.  The pool is a RING, so device memory does not grow with the video: row 0
    holds frame 0 (slot 0 of every window, video_depth.py:363), window j's new frames go to group j % 3 (32 rows each) and
    stay valid for window j + 1, which re-reads up to nine of them (its overlap / key-frame slots); group j % 3 is
    overwritten by window j + 3's upload, which the driver orders after window j + 1's forward.
    Returns ([[(first_id, count, first_row)] per window], [[row of each slot] per window], pool rows)."""
    runs_all, rows_all = [], []
    prev = {}          # id -> row of the frames uploaded for the previous window
    have0 = False
    for j, row in enumerate(idx_rows):
        ids = [int(i) for i in row]
        need = sorted(set(ids))
        cur, runs = {}, []
        base, used = 1 + (j % POOL_GROUPS) * INFER_LEN, 0
        for i in need:
            if (i == 0 and have0) or i in prev:
                continue
            if i == 0:
                r = 0
                have0 = True
            else:
                r = base + used
                used += 1
                cur[i] = r
            if runs and runs[-1][0] + runs[-1][1] == i and runs[-1][2] + runs[-1][1] == r:
                runs[-1][1] += 1
            else:
                runs.append([i, 1, r])
        lookup = dict(prev)
        lookup.update(cur)
        lookup[0] = 0
        rows_all.append([lookup[i] for i in ids])
        runs_all.append([tuple(r) for r in runs])
        prev = cur
    return runs_all, rows_all, 1 + POOL_GROUPS * INFER_LEN


class WindowStitcher:
    """One round of the sharded stitching pass: exchange, (scale, shift) recurrence, finalisation of this rank's window.

    Pure orchestration: the arithmetic goes through `ops` (the C-ABI kernels in the product, `CudaStitchOps`; the CPU
    tests drive the same orchestration with an oracle-backed ops object over gloo), buffers through `alloc`, the exchange
    through `all_gather(dst[world, 11, ...], src[11, ...])` (None for a single rank).
    """

    def __init__(self, ops, n_frames, world, rank, frame_shape, alloc, all_gather=None):
        self.ops, self.n, self.world, self.rank = ops, int(n_frames), int(world), int(rank)
        self.n_win = (self.n + FRAME_STEP - 1) // FRAME_STEP
        self.rounds = (self.n_win + self.world - 1) // self.world
        self.all_gather = all_gather if self.world > 1 else None
        nx = len(XCHG_SLOTS)
        fs = tuple(frame_shape)
        self.xchg = [alloc((self.world, nx) + fs), alloc((self.world, nx) + fs)]   # round j -> buffer j & 1
        self.send = [alloc((nx,) + fs), alloc((nx,) + fs)]
        self.ref = alloc((ALIGN_LEN,) + fs)      # [depth of video frame 0, previous window's aligned slot 12]
        self.st = alloc((self.n_win, 2))         # (scale, shift) of every window; row 0 is never read (identity)
        self.exchange_bytes_per_round = 0 if self.all_gather is None else nx * int(np.prod(fs)) * 4 * (self.world - 1)

    def send_buffer(self, j):
        """Where the owner of this rank's window of round j puts the window's raw XCHG_SLOTS frames before `round(j)`."""
        return self.send[j & 1]

    def exchange(self, j):
        if self.all_gather is not None:
            self.all_gather(self.xchg[j & 1], self.send[j & 1])
        else:   # a single rank: the send buffer is free again for round j + 2 while round j + 1 still reads this round's tail
            self.ops.align(self.send[j & 1], None, self.xchg[j & 1][0])

    def round(self, j, raw, out):
        """raw: [32, ...] raw depth of this rank's window of round j (None if it has none).  Writes the window's final
        frames to out[0:count] and returns (first_frame, count) with count clipped to the video length, or None."""
        ops, X = self.ops, self.xchg[j & 1]
        first = j * self.world
        for r in range(self.world):   # the scalar recurrence, replayed identically on every rank
            k = first + r
            if k >= self.n_win:
                break
            if k == 0:
                ops.align(X[0][0:1], None, self.ref[0:1])
                ops.align(X[0][XCHG_REF:XCHG_REF + 1], None, self.ref[1:2])
            else:
                ops.fit(X[r][XCHG_FIT], self.ref, self.st[k])
                ops.align(X[r][XCHG_REF:XCHG_REF + 1], self.st[k], self.ref[1:2])
        k = first + self.rank
        if k >= self.n_win or raw is None:
            return None
        lo, _ = final_range(k, self.n_win)
        if k == 0:
            cnt = INFER_LEN - INTERP_LEN
            ops.align(raw[0:cnt], None, out[0:cnt])   # window 0 is used as it is (video_depth.py:382-387)
        else:
            prev = X[self.rank - 1][XCHG_TAIL] if self.rank > 0 else self.xchg[(j & 1) ^ 1][self.world - 1][XCHG_TAIL]
            st_prev = self.st[k - 1] if k - 1 > 0 else None
            ops.blend(prev, st_prev, raw[ALIGN_LEN:OVERLAP], self.st[k], out[0:INTERP_LEN])
            ops.align(raw[OVERLAP:INFER_LEN - INTERP_LEN], self.st[k], out[INTERP_LEN:FRAME_STEP])
            cnt = FRAME_STEP
        if k == self.n_win - 1:
            ops.align(raw[INFER_LEN - INTERP_LEN:INFER_LEN], self.st[k] if k > 0 else None, out[cnt:cnt + INTERP_LEN])
            cnt += INTERP_LEN
        return lo, max(0, min(lo + cnt, self.n) - lo)


class CudaStitchOps:
    """The stitching arithmetic on the device, through the C ABI, on the current stream."""

    def __init__(self, lib, frame_elems, device):
        self.lib, self.fe = lib, int(frame_elems)
        n = int(lib.vda_scale_shift_det_workspace_doubles(ALIGN_LEN * self.fe))
        self.partial = torch.empty((n,), dtype=torch.float64, device=device)

    def fit(self, pred, ref, st):
        L.check(self.lib.vda_scale_shift_det(pred.data_ptr(), ref.data_ptr(), pred.shape[0] * self.fe, self.partial.data_ptr(),
                                             st.data_ptr(), L.stream_ptr()), "vda_scale_shift_det")

    def align(self, src, st, dst):
        count = src.shape[0] * self.fe
        if st is None:
            L.check(self.lib.vda_copy_device(src.data_ptr(), dst.data_ptr(), count * 4, L.stream_ptr()), "vda_copy_device")
        else:
            L.check(self.lib.vda_scale_clamp(src.data_ptr(), dst.data_ptr(), st.data_ptr(), count, L.stream_ptr()), "vda_scale_clamp")

    def blend(self, pre, st_pre, post, st_post, dst):
        L.check(self.lib.vda_blend_raw(pre.data_ptr(), L._ptr(st_pre), post.data_ptr(), st_post.data_ptr(), dst.data_ptr(),
                                       pre.shape[0], self.fe, L.stream_ptr()), "vda_blend_raw")


def open_shared_buffer(nbytes, group):
    """A host buffer of `nbytes` mapped by every rank of `group` (one node): anonymous shared memory created by group rank 0
    (memfd: not bounded by the size of the /dev/shm mount) and opened by the others through /proc.  Returns the mmap."""
    dist = torch.distributed
    rank = dist.get_rank(group)
    src = dist.get_global_rank(group, 0) if group is not None else 0
    fd = -1
    name = [None]
    if rank == 0:
        fd = os.memfd_create("vda_b200_result")
        os.ftruncate(fd, nbytes)
        name = [f"/proc/{os.getpid()}/fd/{fd}"]
    dist.broadcast_object_list(name, src=src, group=group)
    if rank != 0:
        fd = os.open(name[0], os.O_RDWR)
    mm = mmap.mmap(fd, nbytes)
    dist.barrier(group)   # everybody holds a mapping before the creator's descriptor goes away
    os.close(fd)
    return mm


class WindowedDepthDriver:
    def __init__(self, model, device="cuda", fp32=False, skip_tmp_block=False, upload_chunk=64, group=None):
        """group: None = this process alone (the reference's behaviour, whatever process group may exist); a
        torch.distributed process group (e.g. `torch.distributed.group.WORLD`) = shard the windows over its ranks."""
        self.model = model
        self.device = torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        L.require_device(self.device)
        self.fp32 = fp32
        self.skip_tmp_block = skip_tmp_block
        self.upload_chunk = upload_chunk
        self.group = group
        self.lib = L.load()
        self.last_scale_shift = None
        self.last_stage_s = None
        self.last_frame_ranges = None

    def _world(self):
        if self.group is None:
            return 1, 0
        dist = torch.distributed
        return dist.get_world_size(self.group), dist.get_rank(self.group)

    # ------------------------------------------------------------------ upload + pre-processing
    def _to_device_u8(self, frames, lo, hi):
        """Source frames [lo, hi) as a device uint8 tensor [hi-lo, H, W, 3] (no copy when they already are one)."""
        if isinstance(frames, torch.Tensor):
            if frames.dtype != torch.uint8:
                raise TypeError("frames must be uint8 [N,H,W,3]")
            part = frames[lo:hi]
            return part if part.device == self.device and part.is_contiguous() else part.contiguous().to(self.device, non_blocking=True)
        if isinstance(frames, np.ndarray) and frames.flags.c_contiguous:
            host = frames[lo:hi]
        else:
            host = np.ascontiguousarray(np.stack([np.asarray(frames[i]) for i in range(lo, hi)]))
        if host.dtype != np.uint8:
            raise TypeError("frames must be uint8 [N,H,W,3]")
        return torch.from_numpy(host).to(self.device, non_blocking=True)

    def _upload_preprocess(self, frames, lo, hi, pre, nh, nw, row0=None):
        """Source frames [lo, hi) -> pre[row0 : row0 + hi - lo] (util/transform.py:109-158 + video_depth.py:361; once per
        source frame).  row0 defaults to lo."""
        fh, fw = frames[0].shape[:2]
        st = L.stream_ptr()
        row0 = lo if row0 is None else row0
        for c0 in range(lo, hi, self.upload_chunk):
            c1 = min(hi, c0 + self.upload_chunk)
            dev = self._to_device_u8(frames, c0, c1)
            L.check(self.lib.vda_preprocess_u8(dev.data_ptr(), pre[row0 + c0 - lo].data_ptr(), c1 - c0, fh, fw, nh, nw, st),
                    "vda_preprocess_u8")
            del dev

    # ------------------------------------------------------------------ whole job
    def run(self, frames, input_size=518, return_device=False, fresh=True):
        """frames: uint8 [N, H, W, 3] -- a numpy array (page-locked memory makes the uploads asynchronous), any object with
        `.shape[0]` and `frames[i]` (what the reference requires, video_depth.py:350), or a torch uint8 tensor (on this
        device: nothing is uploaded).

        Returns the aligned depth [N, H, W] fp32:
          * return_device: this rank's final frames as a device tensor -- the whole stack in order for a single rank; for a
            sharded run the frames of `self.last_frame_ranges` = [(first_frame, count, row)] (nothing crosses PCIe);
          * single rank: a numpy array -- fresh (reference semantics) or, with fresh=False, a view of the driver's cached
            page-locked buffer that the next call overwrites;
          * sharded: group rank 0 gets a numpy view of the host buffer shared by the ranks (valid until the next sharded call
            on this model), the other ranks get None."""
        n = int(frames.shape[0])
        fh, fw = (int(v) for v in frames[0].shape[:2])
        isz = effective_input_size(fh, fw, input_size)
        nh, nw = network_size(fh, fw, isz)
        world, rank = self._world()
        with torch.cuda.device(self.device), torch.no_grad():
            return self._run_pipeline(frames, n, fh, fw, nh, nw, world, rank, return_device, fresh)

    def _run_pipeline(self, frames, n, fh, fw, nh, nw, world, rank, return_device, fresh):
        lib, dev = self.lib, self.device
        dist = torch.distributed
        t_begin = time.perf_counter()
        idx = window_index_map(n)
        n_win = idx.shape[0]
        mine = rank_windows(n_win, world, rank)
        rounds = (n_win + world - 1) // world
        fe = fh * fw
        plan = self.model.plan(1, INFER_LEN, nh, nw, dev, not self.fp32, self.skip_tmp_block)

        main = torch.cuda.current_stream()
        up, side = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        start = torch.cuda.Event()
        start.record(main)
        up.wait_event(start)
        side.wait_event(start)

        # ---- the rank's ring of pre-processed frames and the per-window pointer rows into it
        runs, slot_rows, pool_rows = upload_runs([idx[k] for k in mine])
        pre = torch.empty((pool_rows, 3, nh, nw), dtype=torch.float32, device=dev)
        fbytes = 3 * nh * nw * 4
        if mine:
            table = torch.tensor([[pre.data_ptr() + r * fbytes for r in rows] for rows in slot_rows], dtype=torch.int64)
            table = table.to(dev, non_blocking=True)

        def alloc(shape):
            return torch.empty(shape, dtype=torch.float32, device=dev)

        gather = None
        if world > 1:
            def gather(dst, src):
                dist.all_gather_into_tensor(dst.flatten(0, 1), src, group=self.group)   # rank r's 11 frames -> dst[r]
        ops = CudaStitchOps(lib, fe, dev)
        stitcher = WindowStitcher(ops, n, world, rank, (fh, fw), alloc, gather)
        stitcher.st.zero_()
        stitcher.st[:, 0] = 1.0
        raw = alloc((2, INFER_LEN, fh, fw))                       # round j -> slot j & 1
        my_frames = sum(max(0, min(final_range(k, n_win)[1], n) - final_range(k, n_win)[0]) for k in mine)
        # this rank's final frames: all of them, in processing order, when they stay on the device; otherwise one window's
        # worth that the stitch and its device->host copy (same stream) pass through
        out = alloc(((max(1, my_frames) if return_device else 0) + INFER_LEN, fh, fw))

        # ---- where the result goes
        host = host_np = result = worker = jobs = None
        if not return_device:
            if world > 1:
                host = self._shared_result(n * fe * 4).view(torch.float32)[:n * fe].view(n, fh, fw)
            else:
                host = self._pinned_staging(n * fe * 4).view(torch.float32)[:n * fe].view(n, fh, fw)
            host_np = host.numpy()
            if world == 1 and fresh:
                # device -> page-locked staging -> the caller's fresh pageable array, the second hop on a worker thread
                # while the GPU keeps computing
                result = np.empty((n, fh, fw), dtype=np.float32)
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

        fwd_done = []   # per round: the event after its forward (None when the rank has no window in it)

        def upload(j):
            with torch.cuda.stream(up):
                if j >= 2 and fwd_done[j - 2] is not None:
                    up.wait_event(fwd_done[j - 2])   # group j % 3 was last read by windows j - 3 and j - 2
                for (first, count, row0) in runs[j]:
                    self._upload_preprocess(frames, first, first + count, pre, nh, nw, row0=row0)
                ev = torch.cuda.Event()
                ev.record(up)
            return ev

        timed = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
        ev_round = [None] * rounds
        marks = []       # (fwd0, fwd1, x0, x1, s1, d1) per round
        ranges = []
        pos = 0
        ev_up = upload(0) if mine else None
        try:
            for j in range(rounds):
                k = mine[j] if j < len(mine) else None
                m = {}
                if k is not None:
                    main.wait_event(ev_up)
                    if j >= 2:
                        main.wait_event(ev_round[j - 2])   # raw / send slot j & 1 was last read by round j - 2's stitch
                    s = main.cuda_stream
                    m["f0"] = timed()
                    m["f0"].record(main)
                    L.check(lib.vda_copy_frames(table.data_ptr() + j * INFER_LEN * 8, plan.src_ptrs.data_ptr(), None, None, 1,
                                                INFER_LEN * 8, s), "vda_copy_frames")
                    plan.launch()
                    cur = raw[j & 1]
                    L.check(lib.vda_resize_depth(plan.depth.data_ptr(), L.F32, cur.data_ptr(), INFER_LEN, nh, nw, fh, fw, s),
                            "vda_resize_depth")
                    send = stitcher.send_buffer(j)
                    # the frames the other ranks need, resized straight into the exchange buffer (0, 1 | 12 | 24..31)
                    for dst0, slot0, cnt in ((0, 0, ALIGN_LEN), (XCHG_REF, KEYFRAMES[1], 1),
                                             (XCHG_REF + 1, INFER_LEN - INTERP_LEN, INTERP_LEN)):
                        L.check(lib.vda_resize_depth(plan.depth[slot0].data_ptr(), L.F32, send[dst0].data_ptr(), cnt, nh, nw, fh,
                                                     fw, s), "vda_resize_depth")
                    m["f1"] = timed()
                    m["f1"].record(main)
                    fwd_done.append(m["f1"])
                    side.wait_event(m["f1"])
                    if j + 1 < len(mine):
                        ev_up = upload(j + 1)
                else:
                    fwd_done.append(None)
                with torch.cuda.stream(side):
                    m["x0"] = timed()
                    m["x0"].record(side)
                    stitcher.exchange(j)
                    m["x1"] = timed()
                    m["x1"].record(side)
                    done = stitcher.round(j, raw[j & 1] if k is not None else None, out[pos:] if return_device else out)
                    m["s1"] = timed()
                    m["s1"].record(side)
                    if done is not None and done[1] > 0:
                        lo, cnt = done
                        ranges.append((lo, cnt, pos))
                        if host is not None:
                            host[lo:lo + cnt].copy_(out[:cnt], non_blocking=True)
                            if jobs is not None:
                                landed = torch.cuda.Event()
                                landed.record(side)
                                jobs.put((landed, lo, lo + cnt))
                        pos += cnt
                    m["d1"] = timed()
                    m["d1"].record(side)
                    ev_round[j] = m["d1"]
                marks.append(m)
        finally:
            # give the plan its own input frames back (`forward()` reads x_in through the same pointer array)
            L.check(lib.vda_copy_frames(plan.src_ptrs_default.data_ptr(), plan.src_ptrs.data_ptr(), None, None, 1, INFER_LEN * 8,
                                        main.cuda_stream), "vda_copy_frames")
        main.wait_stream(side)
        main.wait_stream(up)
        self.last_scale_shift = stitcher.st
        self.last_frame_ranges = ranges
        if return_device:
            self._stage_times(marks, t_begin, stitcher, sync=False)
            return out[:n] if world == 1 else out[:pos]
        if jobs is not None:
            jobs.put(None)
            worker.join()
        main.synchronize()
        if world > 1:
            dist.barrier(self.group)   # every rank's frames have landed in the shared buffer
        self._stage_times(marks, t_begin, stitcher, sync=True)
        if world > 1:
            return host_np if rank == 0 else None
        return result if result is not None else host_np

    def _stage_times(self, marks, t_begin, stitcher, sync):
        """Device-side stage times of this call from CUDA events on the launching streams (summed over the rank's rounds)."""
        info = {"rounds": len(marks), "exchange_bytes_per_round": stitcher.exchange_bytes_per_round}
        if sync:
            f = x = s = d = 0.0
            for m in marks:
                if "f0" in m:
                    f += m["f0"].elapsed_time(m["f1"])
                x += m["x0"].elapsed_time(m["x1"])
                s += m["x1"].elapsed_time(m["s1"])
                d += m["s1"].elapsed_time(m["d1"])
            info.update(forward_s=f / 1e3, exchange_s=x / 1e3, stitch_s=s / 1e3, d2h_s=d / 1e3, total_s=time.perf_counter() - t_begin)
        self.last_stage_s = info

    # ------------------------------------------------------------------ host buffers
    def _download(self, depth, chunk_frames=32, workers=4):
        """Device [n, H, W] fp32 -> a fresh pageable numpy array (used by the streaming mode).  A plain `.cpu()` into
        pageable memory moves ~2 GB/s (the runtime stages through small internal buffers and the page faults of the new
        array land on one thread); here the stack goes to the cached page-locked staging buffer in chunks on a side stream
        at PCIe rate while worker threads copy finished chunks out (numpy releases the GIL for the memcpy, the first-touch
        faults spread over the threads)."""
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

    def _shared_result(self, nbytes):
        """The node-wide result buffer of a sharded run: shared memory mapped by every rank of the group and page-locked in
        each process, so that every rank's device->host copies are asynchronous DMA over its own PCIe link.  Cached on the
        model (creating and page-locking gigabytes costs seconds) and grown on demand; all ranks call this together."""
        hit = getattr(self.model, "_vda_shared_result", None)
        if hit is None or hit.tensor.numel() < nbytes:
            if hit is not None:
                hit.release()
            hit = SharedHostBuffer(nbytes, self.group, self.lib)
            self.model._vda_shared_result = hit
        return hit.tensor


class SharedHostBuffer:
    """Shared memory mapped by every rank of `group`, page-locked in this process (released explicitly or on collection:
    a mapping must never go away while the CUDA runtime still holds it registered)."""

    def __init__(self, nbytes, group, lib):
        self.lib = lib
        self.mm = open_shared_buffer(nbytes, group)
        self.tensor = torch.frombuffer(self.mm, dtype=torch.uint8)
        self.registered = lib.vda_host_register(self.tensor.data_ptr(), nbytes) == 0
        if not self.registered:
            warnings.warn("video_depth_anything_b200: page-locking the shared result buffer failed ("
                          f"{lib.vda_last_error().decode(errors='replace')}); device->host copies into it will be staged by the "
                          "runtime", RuntimeWarning)

    def release(self):
        if self.registered:
            torch.cuda.synchronize()
            self.lib.vda_host_unregister(self.tensor.data_ptr())
            self.registered = False

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass

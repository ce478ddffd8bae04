"""The N>1 host path of the window-sharded `infer_video_depth` on CPU: round-robin window
ownership, the per-round all-gather of 11 frames per window, the redundantly replayed (scale, shift) recurrence, the
finalisation of each rank's own windows and the shared host buffer -- `driver.WindowStitcher` / `open_shared_buffer`, the
very orchestration the CUDA driver runs, here over gloo with an ops object backed by the oracle's numpy arithmetic
(test infrastructure: the product only ever plugs in `CudaStitchOps`).  Every rank fabricates the raw depth of ITS
windows (a deterministic function of window and source frame); the assembled result must equal, bit for bit, the
oracle's single-process stitching pass (reference video_depth.py:379-417) over all windows."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from video_depth_anything_b200.driver import (FRAME_STEP, INFER_LEN, XCHG_SLOTS, WindowStitcher, final_range, open_shared_buffer,
                                              rank_windows, upload_runs, window_index_map)


class OracleOps:
    """`ops` of WindowStitcher on CPU torch tensors with the oracle's arithmetic (oracle/driver.py)."""

    def fit(self, pred, ref, st):
        from oracle import driver as odrv
        p, r = pred.numpy(), ref.numpy()
        s, t = odrv.compute_scale_and_shift(np.concatenate(list(p)), np.concatenate(list(r)), np.concatenate(list(np.ones_like(r) == 1)))
        st[0], st[1] = float(np.float32(s)), float(np.float32(t))

    def align(self, src, st, dst):
        if st is None:
            dst.copy_(src)
            return
        s, t = np.float32(st[0].item()), np.float32(st[1].item())
        v = src.numpy() * s + t
        v[v < 0] = 0
        dst.copy_(torch.from_numpy(v))

    def blend(self, pre, st_pre, post, st_post, dst):
        from oracle import driver as odrv
        a, b = torch.empty_like(pre), torch.empty_like(post)
        self.align(pre, st_pre, a)
        self.align(post, st_post, b)
        out = odrv.blend_frames(list(a.numpy()), list(b.numpy()))
        dst.copy_(torch.from_numpy(np.stack(out).astype(np.float32)))


def _fake_window(wid, row, h, w):
    """[32, h, w] raw depth of window `wid`: a different affine map of a frame-dependent pattern per window, so that the
    least-squares fit, the clamp and the cross-fade all matter."""
    yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    out = np.empty((INFER_LEN, h, w), np.float32)
    gain, bias = 1.0 + 0.07 * ((wid * 5) % 7), -0.4 + 0.11 * (wid % 4)
    for s, src in enumerate(row):
        base = 1.0 + 0.01 * src + 0.6 * np.sin(0.3 * yy + 0.2 * xx + 0.37 * src)
        out[s] = np.maximum(gain * base + bias + 0.002 * s, 0.0)   # the network's output is non-negative (ReLU)
    return torch.from_numpy(out)


def _run_rank(n_frames, h, w, world, rank, all_gather, result):
    """The loop of WindowedDepthDriver._run_pipeline without the device: returns the (scale, shift) table."""
    idx = window_index_map(n_frames)
    st = WindowStitcher(OracleOps(), n_frames, world, rank, (h, w), lambda shape: torch.zeros(shape, dtype=torch.float32), all_gather)
    mine = rank_windows(idx.shape[0], world, rank)
    out = torch.zeros((INFER_LEN, h, w))
    for j in range(st.rounds):
        raw = None
        if j < len(mine):
            raw = _fake_window(mine[j], idx[mine[j]], h, w)
            st.send_buffer(j).copy_(raw[XCHG_SLOTS])
        st.exchange(j)
        done = st.round(j, raw, out)
        if done is not None and done[1] > 0:
            lo, cnt = done
            result[lo:lo + cnt] = out[:cnt].numpy()
    return st.st.numpy().copy()


def _oracle_stitch(n_frames, h, w):
    from oracle import driver as odrv
    idx = window_index_map(n_frames)
    flat = []
    for k in range(idx.shape[0]):
        flat += list(_fake_window(k, idx[k], h, w).numpy())
    return odrv.stitch(flat, n_frames)


def test_schedule_covers_every_frame_once():
    for n in (1, 5, 22, 23, 32, 44, 45, 70, 100, 1024, 2048):
        n_win = window_index_map(n).shape[0]
        assert n_win == (n + FRAME_STEP - 1) // FRAME_STEP
        pos = 0
        for k in range(n_win):
            lo, hi = final_range(k, n_win)
            assert lo == pos
            pos = hi
        assert pos >= n
        for world in (1, 2, 3, 8):
            owned = sorted(k for r in range(world) for k in rank_windows(n_win, world, r))
            assert owned == list(range(n_win))
    assert rank_windows(94, 8, 7) == list(range(7, 94, 8)) and len(rank_windows(94, 8, 5)) == 12
    # the ring of pre-processed frames: every slot of every window resolves to the row its frame was last uploaded to, a
    # group is only rewritten three windows later, and frame 0 keeps row 0
    for world, rank in ((1, 0), (2, 1), (8, 3)):
        idx = window_index_map(2048)
        wins = [idx[k] for k in rank_windows(94, world, rank)]
        runs, slot_rows, pool = upload_runs(wins)
        assert pool == 97 and len(runs) == len(slot_rows) == len(wins)
        content = {}   # row -> frame id, replayed in upload order
        for j, (win, rr, sr) in enumerate(zip(wins, runs, slot_rows)):
            for first, count, row0 in rr:
                for t in range(count):
                    row = row0 + t
                    assert row == 0 or 1 + (j % 3) * 32 <= row < 1 + (j % 3) * 32 + 32
                    content[row] = first + t
            assert [content[r] for r in sr] == [int(i) for i in win]
            assert max(len(rr), 1) <= 4
        # a single rank uploads every frame exactly once (22 new frames per window after the first)
        if world == 1:
            assert sum(c for rr in runs for (_, c, _) in rr) == 2048


@pytest.mark.parametrize("n_frames", [5, 22, 23, 45, 100])
def test_single_rank_rounds_equal_the_oracle_stitch(n_frames):
    h, w = 6, 10
    result = np.zeros((n_frames, h, w), np.float32)
    st = _run_rank(n_frames, h, w, 1, 0, None, result)
    ref, st_ref = _oracle_stitch(n_frames, h, w)
    assert np.array_equal(result, ref)
    assert np.array_equal(st[1:], np.asarray(st_ref, np.float32)[1:])


def _worker(rank, world, port, n_frames, h, w, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mm = open_shared_buffer(n_frames * h * w * 4, dist.group.WORLD)
        result = np.frombuffer(mm, dtype=np.float32).reshape(n_frames, h, w)

        def gather(dst, src):
            dist.all_gather_into_tensor(dst.flatten(0, 1), src)   # the call the CUDA driver makes

        st = _run_rank(n_frames, h, w, world, rank, gather, result)
        dist.barrier()
        q.put((rank, st, result.copy() if rank == 0 else None))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,n_frames", [(2, 100), (3, 150)])   # 5 windows over 2 ranks, 7 over 3: ragged last rounds
def test_sharded_rounds_over_gloo_equal_the_single_process_stitch(world, n_frames):
    h, w = 6, 10
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, h, w, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    ref, st_ref = _oracle_stitch(n_frames, h, w)
    sts = {r: st for r, st, _ in got}
    full = [res for r, _, res in got if r == 0][0]
    for r in range(world):   # every rank replayed the same recurrence
        assert np.array_equal(sts[r], sts[0])
    assert np.array_equal(sts[0][1:], np.asarray(st_ref, np.float32)[1:])
    assert np.array_equal(full, ref) and full.shape == (n_frames, h, w)

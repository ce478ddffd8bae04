"""Window sharding + gather on 2 ranks over gloo (CPU): the N>1 host path of `infer_video_depth`.
Each rank fabricates the raw depth blocks of ITS windows (a deterministic function of the window's source-frame
indices), rank 0 gathers and must hold exactly the stack a single rank would have produced, and the oracle's stitching
pass over it must equal the single-process result."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from video_depth_anything_b200.driver import INFER_LEN, gather_window_blocks, shard_windows, window_index_map


def _fake_depth(idx_rows, h, w):
    """[n_win*32, h, w]: slot value depends only on (window position in the video, source frame index)."""
    yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    out = np.empty((len(idx_rows) * INFER_LEN, h, w), np.float32)
    for k, (wid, row) in enumerate(idx_rows):
        for s, src in enumerate(row):
            out[k * INFER_LEN + s] = 1.0 + 0.01 * src + 0.001 * wid + 0.05 * np.sin(0.3 * yy + 0.2 * xx + src)
    return torch.from_numpy(out)


def _worker(rank, world, port, n_frames, h, w, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    idx = window_index_map(n_frames)
    first, count = shard_windows(idx.shape[0], world)[rank]
    mine = [(first + k, idx[first + k]) for k in range(count)]
    raw = _fake_depth(mine, h, w) if count else torch.empty((0, h, w))
    full = gather_window_blocks(raw, idx.shape[0], world, rank)
    if rank == 0:
        q.put(full.numpy())
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_windows_partition():
    for n_win in (1, 2, 5, 47, 94):
        for world in (1, 2, 4, 8):
            sh = shard_windows(n_win, world)
            assert len(sh) == world and sum(c for _, c in sh) == n_win
            pos = 0
            for lo, c in sh:
                assert c == 0 or lo == pos
                pos += c
    assert [c for _, c in shard_windows(47, 8)] == [6, 6, 6, 6, 6, 6, 6, 5]


def test_two_rank_gather_matches_single_rank_and_stitches():
    from oracle import driver as odrv
    n_frames, h, w, world = 100, 12, 16, 2   # 5 windows -> 3 + 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, h, w, q)) for r in range(world)]
    for p in procs:
        p.start()
    full = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    idx = window_index_map(n_frames)
    single = _fake_depth(list(enumerate(idx)), h, w).numpy()
    assert np.array_equal(full, single)
    a, _ = odrv.stitch([d for d in full], n_frames)
    b, _ = odrv.stitch([d for d in single], n_frames)
    assert np.array_equal(np.asarray(a), np.asarray(b)) and np.asarray(a).shape == (n_frames, h, w)

"""Video ingest / egress: oracle vs the goldens made from the real reference
(scripts/make_goldens.py --io), the product's host-side ingest vs the same goldens, and -- on the GPU -- the device
visualisation kernels vs the reference's pictures (bit-exact: uint8 output of fp32 arithmetic)."""
import hashlib
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")
CLIP = os.path.join(GOLD, "clip_48x64.mp4")
READ_CASES = [(-1, -1, -1), (10, -1, -1), (-1, 10, -1), (17, 12, 40), (-1, 100, -1), (5, -1, 32)]
DEPTH_CASES = [("odd", 11, 5, 37, 41, False), ("vec", 12, 3, 32, 48, True)]
MODES = (("grey", dict(grayscale=True)), ("inferno", dict()), ("spectral", dict(spectral=True)),
         ("spectral_grey", dict(grayscale=True, spectral=True)))


def _gold():
    return np.load(os.path.join(GOLD, "video_io.npz"))


def _check_read(fn):
    g = _gold()
    for k, (plen, tfps, mres) in enumerate(READ_CASES):
        assert tuple(g[f"read{k}_args"]) == (plen, tfps, mres)
        frames, fps = fn(CLIP, plen, tfps, mres)
        want = g[f"read{k}_shape_fps"]
        assert frames.dtype == np.uint8 and list(frames.shape) == [int(v) for v in want[:4]], (k, frames.shape)
        assert fps == want[4]
        assert np.array_equal(frames[0], g[f"read{k}_first"]), k
        sha = np.frombuffer(hashlib.sha256(np.ascontiguousarray(frames).tobytes()).digest(), dtype=np.uint8)
        assert np.array_equal(sha, g[f"read{k}_sha"]), k


def test_oracle_read_video_frames_matches_reference_goldens():
    from oracle import video_io as oio
    _check_read(oio.read_video_frames)
    assert oio.frame_selection(24, 30.0, 17, 12) == ([0, 2, 4, 6, 8, 10, 12, 14, 16], 12)
    assert oio.frame_selection(24, 30.0, -1, 100)[0] == list(range(24))
    assert oio.shrunk_size(48, 64, 40) == (30, 40) and oio.shrunk_size(48, 64, 64) == (48, 64)


def test_product_read_video_frames_matches_reference_goldens():
    from video_depth_anything_b200 import video_io as vio
    _check_read(vio.read_video_frames)
    frames, _ = vio.read_video_frames(CLIP, -1, -1, -1)
    assert frames.flags.c_contiguous and frames.base is not None  # one block, handed out as a view
    with pytest.raises(ValueError, match="need at least one array to stack"):   # np.stack([]) in the reference
        vio.read_video_frames(os.path.join(GOLD, "no_such_file.mp4"), -1, 24, -1)
    with pytest.raises(ZeroDivisionError):   # original_fps / fps with a container that reports 0 fps (dc_utils.py:52)
        vio.read_video_frames(os.path.join(GOLD, "no_such_file.mp4"), -1, -1, -1)


def test_frame_store_grows_past_its_estimate():
    from video_depth_anything_b200.video_io import FrameStore
    st = FrameStore(4, 6, 2)
    for i in range(7):
        st.slot()[...] = i
    out = st.frames()
    assert out.shape == (7, 4, 6, 3) and [int(out[i, 0, 0, 0]) for i in range(7)] == list(range(7))


def test_oracle_depth_pictures_match_reference_goldens():
    from oracle import video_io as oio
    from oracle.inputs import synthetic_colour_tables, synthetic_depths
    g = _gold()
    inferno, spectral = synthetic_colour_tables(0)
    for name, seed, n, h, w, signed in DEPTH_CASES:
        d = synthetic_depths(seed, n, h, w, signed)
        for mode, kw in MODES:
            assert np.array_equal(oio.depth_pictures(d, inferno=inferno, spectral_table=spectral, **kw), g[f"depth_{name}_{mode}"])


def test_colour_table_rules():
    from video_depth_anything_b200 import video_io as vio
    from oracle.inputs import synthetic_colour_tables
    inferno, _ = synthetic_colour_tables(0)
    assert vio._colour_table(True, False, None) is None
    t = vio._colour_table(False, False, inferno)
    assert t.dtype == np.uint8 and np.array_equal(t, (inferno * 255).astype(np.uint8))
    assert vio._colour_table(True, True, t) is not None   # spectral wins over grayscale (dc_utils.py:78-81)
    with pytest.raises(ValueError):
        vio._colour_table(False, False, inferno[:100])
    try:
        import matplotlib  # noqa: F401
    except ImportError:
        with pytest.raises(ImportError):   # the reference fails at `import matplotlib.cm` as well
            vio._colour_table(False, False, None)


@pytest.mark.gpu
def test_depth_pictures_on_device_bit_exact():
    from video_depth_anything_b200 import video_io as vio
    from oracle.inputs import synthetic_colour_tables, synthetic_depths
    g = _gold()
    inferno, spectral = synthetic_colour_tables(0)
    for name, seed, n, h, w, signed in DEPTH_CASES:
        d = synthetic_depths(seed, n, h, w, signed)
        for mode, kw in MODES:
            table = spectral[:, :3] if kw.get("spectral") else inferno
            for src in (d, torch.from_numpy(d).cuda()):          # host array (the reference API) or device tensor
                for chunk in (64, 2):                            # one chunk / several chunks through both staging buffers
                    got = vio.depth_pictures(src, colormap=table, device="cuda:0", chunk_frames=chunk, **kw)
                    assert got.dtype == np.uint8 and np.array_equal(got, g[f"depth_{name}_{mode}"]), (name, mode, chunk)


@pytest.mark.gpu
def test_depth_pictures_full_size_properties():
    """518x518 x 64 frames (BASELINE-sized maps): bit-exact vs the oracle on the same input, min -> 0 and max -> 255."""
    from video_depth_anything_b200 import video_io as vio
    from oracle import video_io as oio
    from oracle.inputs import synthetic_colour_tables, synthetic_depths
    inferno, spectral = synthetic_colour_tables(0)
    d = synthetic_depths(21, 64, 518, 518)
    got = vio.depth_pictures(d, grayscale=True, device="cuda:0")
    assert np.array_equal(got, oio.depth_pictures(d, grayscale=True))
    assert got.min() == 0 and got.max() == 255
    assert got.reshape(-1)[np.argmax(d)] == 255 and got.reshape(-1)[np.argmin(d)] == 0
    col = vio.depth_pictures(d, colormap=inferno, device="cuda:0")
    assert np.array_equal(col, (inferno * 255).astype(np.uint8)[got])
    const = vio.depth_pictures(np.full((2, 8, 8), 3.0, np.float32), grayscale=True, device="cuda:0")
    assert const.max() == 0   # max == min: numpy casts a NaN, 0 on x86


@pytest.mark.gpu
def test_save_video_round_trip(tmp_path):
    """save_video(is_depths) -> file -> read_video_frames: the container holds N frames of the right size; pinned ingest
    feeds the windowed driver directly."""
    from video_depth_anything_b200 import video_io as vio
    from oracle.inputs import synthetic_depths
    d = synthetic_depths(5, 6, 48, 64)
    path = str(tmp_path / "vis.mp4")
    vio.save_video(d, path, fps=12, is_depths=True, grayscale=True, device="cuda:0")
    frames, fps = vio.read_video_frames(path, -1, -1, -1)
    assert frames.shape == (6, 48, 64, 3) and fps == 12
    pics = vio.depth_pictures(d, grayscale=True, device="cuda:0")
    assert np.abs(frames[..., 0].astype(np.int32) - pics.astype(np.int32)).mean() < 12   # lossy codec
    assert torch.from_numpy(frames).is_pinned()

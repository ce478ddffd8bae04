"""TEST INFRASTRUCTURE -- CPU restatement of the reference's video ingest / egress arithmetic.
Only tests/, __graft_entry__.smoke() and bench.py's CPU baseline may import this package; the product never does.

Pinned: `scripts/make_goldens.py --io` runs the REAL `utils/dc_utils.py::read_video_frames` and `save_video` (their
module-level `imageio` / `matplotlib.cm` imports satisfied by recording stubs: imageio's writer only receives the finished
pictures, and the colour tables are injected data) and asserts both functions below equal them exactly; the outputs are
the goldens `tests/golden/video_io_*.npz` + `tests/golden/clip_48x64.mp4`.
"""
import numpy as np


def frame_selection(n_decodable: int, original_fps: float, process_length: int, target_fps: float = -1):
    """Indices of the decoded frames `read_video_frames` keeps and the fps it returns (utils/dc_utils.py:50-64):
    fps = original unless target_fps >= 0; stride = max(round(original / fps), 1); decoding stops once `process_length > 0`
    frames have been DECODED (the test is on the running frame count, :56), every stride-th of them is kept."""
    fps = original_fps if target_fps < 0 else target_fps
    stride = max(round(original_fps / fps), 1)
    stop = n_decodable if process_length <= 0 else min(n_decodable, process_length)
    return [i for i in range(stop) if i % stride == 0], fps


def shrunk_size(height: int, width: int, max_res: int):
    """(height, width) after the `max_res` rule of the cv2 branch (utils/dc_utils.py:46-49): no rounding to even there."""
    if max_res > 0 and max(height, width) > max_res:
        scale = max_res / max(height, width)
        return round(height * scale), round(width * scale)
    return height, width


def read_video_frames(video_path, process_length, target_fps=-1, max_res=-1):
    """utils/dc_utils.py:40-68 restated on top of `frame_selection` / `shrunk_size`: decode everything, keep the selected
    frames, BGR -> RGB, shrink with cv2's default (bilinear) interpolation."""
    import cv2
    cap = cv2.VideoCapture(video_path)
    original_fps = cap.get(cv2.CAP_PROP_FPS)
    h, w = int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT)), int(cap.get(cv2.CAP_PROP_FRAME_WIDTH))
    decoded = []
    while cap.isOpened():
        ok, frame = cap.read()
        if not ok:
            break
        decoded.append(frame)
    cap.release()
    keep, fps = frame_selection(len(decoded), original_fps, process_length, target_fps)
    oh, ow = shrunk_size(h, w, max_res)
    out = []
    for i in keep:
        rgb = decoded[i][:, :, ::-1]
        if (oh, ow) != (h, w):
            rgb = cv2.resize(np.ascontiguousarray(rgb), (ow, oh))
        out.append(rgb)
    return np.stack(out, axis=0), fps


def depth_pictures(depths, grayscale=False, spectral=False, inferno=None, spectral_table=None):
    """The pictures `save_video(depths, ..., is_depths=True)` appends to its writer (utils/dc_utils.py:72-82).
    `inferno`: float [256, 3] (matplotlib's `cm.get_cmap("inferno").colors`); `spectral_table`: float [256, 4]
    (`cm.Spectral` evaluated at the integers 0..255, which indexes its 256-entry table)."""
    depths = np.asarray(depths)
    d_min, d_max = depths.min(), depths.max()
    out = []
    for i in range(depths.shape[0]):
        u = ((depths[i] - d_min) / (d_max - d_min) * 255).astype(np.uint8)
        if spectral:
            out.append((np.asarray(spectral_table)[u][:, :, :3] * 255).astype(np.uint8))
        elif grayscale:
            out.append(u)
        else:
            out.append((np.asarray(inferno)[u] * 255).astype(np.uint8))
    return np.stack(out, axis=0)

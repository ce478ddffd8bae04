"""Video ingest / egress either side of the depth path: the reference's
`read_video_frames` and `save_video` (utils/dc_utils.py:19-89, called from run.py:88,157-160) with the same names,
arguments, return values and error behaviour.

What is B200-native here and what is not:
  * ingest: the codec is the host's (`cv2.VideoCapture`; decord, NVDEC and the Video Codec SDK headers are not in this
    image -- the reference falls back to the same cv2 branch, dc_utils.py:40-67).  The decoded RGB frames are written
    straight into ONE page-locked array (`cv2.cvtColor(..., dst=slot)`), so `infer_video_depth` / `infere_single_image`
    upload them by asynchronous DMA under the previous window's kernels instead of through the driver's pageable staging
    copy (which synchronises the stream).
  * egress: the depth visualisation (global min / max, normalisation to uint8, colour table) runs on the device
    (`vda_depth_minmax`, `vda_depth_vis`; csrc/vis.cu) and is bit-identical to the reference's numpy arithmetic; what
    comes back over PCIe is the uint8 picture.  `frames` may be the host array the reference API returns or a device
    tensor.  Encoding is the host's: imageio / libx264 as in the reference when imageio is importable, else
    `cv2.VideoWriter` ('mp4v') -- a different codec, so egress parity is defined on the pictures handed to the encoder.
  * the colour tables are matplotlib data (`cm.get_cmap("inferno").colors`, `cm.Spectral`); matplotlib is not in this image,
    so `save_video` takes the table from matplotlib when it is importable and otherwise needs `colormap=` (float [256, 3] in
    [0, 1] like matplotlib's, or uint8) -- `grayscale=True` needs none.
"""
from __future__ import annotations

import numpy as np
import torch

from . import lib as L

__all__ = ["read_video_frames", "save_video", "depth_pictures", "FrameStore"]


def _pinned_u8(shape):
    """Page-locked uint8 array (numpy view + the torch tensor that owns it); plain memory when there is no CUDA device
    (CPU-side tests of the host logic)."""
    if torch.cuda.is_available():
        t = torch.empty(shape, dtype=torch.uint8, pin_memory=True)
    else:
        t = torch.empty(shape, dtype=torch.uint8)
    return t.numpy(), t


class FrameStore:
    """Growable [n, H, W, 3] uint8 array in page-locked memory; `slot()` hands out the next frame to decode into."""

    def __init__(self, h: int, w: int, capacity: int):
        self.h, self.w = int(h), int(w)
        self.count = 0
        self._arr, self._own = _pinned_u8((max(int(capacity), 1), self.h, self.w, 3))

    def slot(self) -> np.ndarray:
        if self.count == self._arr.shape[0]:  # the container's frame count was an under-estimate
            arr, own = _pinned_u8((2 * self._arr.shape[0], self.h, self.w, 3))
            arr[:self.count] = self._arr
            self._arr, self._own = arr, own
        self.count += 1
        return self._arr[self.count - 1]

    def frames(self) -> np.ndarray:
        return self._arr[:self.count]  # a view: its .base chain keeps the page-locked block alive


def read_video_frames(video_path, process_length, target_fps=-1, max_res=-1):
    """utils/dc_utils.py:19-68 (the cv2 branch :40-66; decord is absent here as in the reference's fallback).
    Returns (uint8 [N, H, W, 3] RGB frames, fps).  Semantics kept: `fps` is the container's unless `target_fps >= 0`; every
    `stride = max(round(original_fps / fps), 1)`-th decoded frame is kept; `process_length > 0` bounds the number of
    DECODED frames (not of kept ones, :56); frames are shrunk with `cv2.resize` (bilinear) when the longer side exceeds
    `max_res > 0`; an unreadable file ends in numpy's "need at least one array to stack" ValueError (:65) -- or in the
    ZeroDivisionError of `original_fps / fps` when the container reports 0 fps and no target is given (:52)."""
    import cv2
    cap = cv2.VideoCapture(video_path)
    src_fps = cap.get(cv2.CAP_PROP_FPS)
    src_h, src_w = int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT)), int(cap.get(cv2.CAP_PROP_FRAME_WIDTH))
    shrink = max_res > 0 and max(src_h, src_w) > max_res
    out_h, out_w = src_h, src_w
    if shrink:
        k = max_res / max(src_h, src_w)
        out_h, out_w = round(src_h * k), round(src_w * k)
    fps = src_fps if target_fps < 0 else target_fps
    stride = max(round(src_fps / fps), 1)
    total = int(cap.get(cv2.CAP_PROP_FRAME_COUNT))
    if process_length > 0:
        total = min(total, process_length) if total > 0 else process_length
    store = FrameStore(out_h, out_w, (max(total, 1) + stride - 1) // stride)
    decoded = 0
    while cap.isOpened():
        ok, bgr = cap.read()
        if not ok or (process_length > 0 and decoded >= process_length):
            break
        if decoded % stride == 0:
            slot = store.slot()
            if shrink:
                cv2.resize(cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB), (out_w, out_h), dst=slot)
            else:
                cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB, dst=slot)
        decoded += 1
    cap.release()
    if store.count == 0:
        raise ValueError("need at least one array to stack")
    return store.frames(), fps


def _colour_table(grayscale, spectral, colormap):
    """uint8 [256, 3] table = what the reference's `(colormap[u] * 255).astype(np.uint8)` /
    `(cm.Spectral(u)[:, :, :3] * 255).astype(np.uint8)` yields for u = 0..255 (dc_utils.py:72,78-81); None for greyscale."""
    if grayscale and not spectral:
        return None
    if colormap is None:
        try:
            import matplotlib.cm as cm
        except ImportError as e:
            raise ImportError("save_video(is_depths=True) without grayscale needs matplotlib's colour tables "
                              "(as the reference does) or an explicit colormap=[256,3] argument") from e
        colormap = cm.Spectral(np.arange(256, dtype=np.uint8))[:, :3] if spectral else np.array(cm.get_cmap("inferno").colors)
    table = np.asarray(colormap)
    if table.shape != (256, 3):
        raise ValueError(f"colormap must be [256, 3], got {table.shape}")
    if table.dtype != np.uint8:
        table = (table * 255).astype(np.uint8)
    return np.ascontiguousarray(table)


def depth_pictures(frames, grayscale=False, spectral=False, colormap=None, device="cuda", chunk_frames=64):
    """The uint8 pictures `save_video(..., is_depths=True)` hands to its encoder (dc_utils.py:72-82): [N, H, W] for
    `grayscale=True`, else [N, H, W, 3].  `frames`: float32 [N, H, W], host ndarray or device tensor.  All per-pixel work
    runs on the device; the result is a host ndarray."""
    lib = L.load()
    table = _colour_table(grayscale, spectral, colormap)
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("depth_pictures runs on a CUDA device only (no CPU path)")
    with torch.cuda.device(dev):
        s = L.stream_ptr()
        if isinstance(frames, torch.Tensor):
            d = frames.to(device=dev, dtype=torch.float32).contiguous()
        else:
            host = np.ascontiguousarray(np.asarray(frames))
            if host.dtype != np.float32:  # the reference's depth stacks are float32; numpy would normalise in the array's dtype
                raise TypeError(f"depth frames must be float32, got {host.dtype}")
            d = torch.empty(host.shape, dtype=torch.float32, device=dev)
            for c0 in range(0, host.shape[0], chunk_frames):
                d[c0:c0 + chunk_frames].copy_(torch.from_numpy(host[c0:c0 + chunk_frames]), non_blocking=True)
        if d.dim() != 3:
            raise ValueError(f"depth frames must be [N, H, W], got {tuple(d.shape)}")
        n, h, w = d.shape
        ch = 1 if table is None else 3
        chunk_frames = max(1, min(int(chunk_frames), n))
        partial = torch.empty(int(lib.vda_depth_minmax_workspace_floats()), dtype=torch.float32, device=dev)
        mm = torch.empty(2, dtype=torch.float32, device=dev)
        L.check(lib.vda_depth_minmax(d.data_ptr(), d.numel(), partial.data_ptr(), mm.data_ptr(), s), "vda_depth_minmax")
        lut = torch.from_numpy(table).to(dev) if table is not None else None
        out = np.empty((n, h, w) if ch == 1 else (n, h, w, 3), dtype=np.uint8)
        # chunks of frames: visualise -> pinned staging (two buffers, so chunk k's copy-out overlaps chunk k+1's kernel)
        stage = [torch.empty((chunk_frames, h, w, ch), dtype=torch.uint8, pin_memory=True) for _ in range(2)]
        pics = [torch.empty((chunk_frames, h, w, ch), dtype=torch.uint8, device=dev) for _ in range(2)]
        inflight = []  # (buffer, event, first frame, last frame) of copies not yet unpacked into `out`

        def drain(buffer=None):
            while inflight and (buffer is None or any(e[0] == buffer for e in inflight)):
                b_, ev, p0, p1 = inflight.pop(0)
                ev.synchronize()
                out[p0:p1] = stage[b_][:p1 - p0].numpy().reshape(out[p0:p1].shape)

        for k, c0 in enumerate(range(0, n, chunk_frames)):
            c1 = min(n, c0 + chunk_frames)
            b = k & 1
            drain(b)  # the staging buffer is about to be reused
            L.check(lib.vda_depth_vis(d[c0].data_ptr(), (c1 - c0) * h * w, mm.data_ptr(), lut.data_ptr() if lut is not None else None,
                                      pics[b].data_ptr(), s), "vda_depth_vis")
            stage[b][:c1 - c0].copy_(pics[b][:c1 - c0], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            inflight.append((b, ev, c0, c1))
        drain()
        return out


def save_video(frames, output_video_path, fps=10, is_depths=False, grayscale=False, spectral=False, colormap=None,
               device="cuda"):
    """utils/dc_utils.py:70-89.  `is_depths=True`: depth maps -> `depth_pictures` (device) -> encoder; otherwise the
    uint8 frames go to the encoder as they are.  The two extra keyword arguments are described in the module docstring."""
    pictures = depth_pictures(frames, grayscale, spectral, colormap, device) if is_depths else np.asarray(frames)
    _encode(pictures, output_video_path, fps)


def _encode(pictures, path, fps):
    try:
        import imageio
    except ImportError:
        imageio = None
    if imageio is not None:  # the reference's writer (dc_utils.py:71)
        writer = imageio.get_writer(path, fps=fps, macro_block_size=1, codec='libx264', ffmpeg_params=['-crf', '18'])
        for i in range(pictures.shape[0]):
            writer.append_data(pictures[i])
        writer.close()
        return
    import cv2
    colour = pictures.ndim == 4
    h, w = pictures.shape[1:3]
    writer = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"mp4v"), float(fps), (w, h), isColor=colour)
    if not writer.isOpened():
        raise IOError(f"cannot open a video writer for {path}")
    for i in range(pictures.shape[0]):
        writer.write(cv2.cvtColor(pictures[i], cv2.COLOR_RGB2BGR) if colour else pictures[i])
    writer.release()

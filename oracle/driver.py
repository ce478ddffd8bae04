"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement (numpy + cv2 4.x `cv2.resize`, the same third-party call the
reference makes at util/transform.py:113) of the sliding-window driver
`VideoDepthAnything.infer_video_depth`
(/root/reference/video_depth_anything/video_depth.py:329-417), its
pre-processing (util/transform.py:51-158) and the alignment math
(/root/reference/utils/util.py:40-74).

The network forward is a parameter (`forward_fn`) so the same driver can sit on
top of the torch oracle (`oracle.model.forward`) on CPU or CUDA.
"""
from __future__ import annotations

import numpy as np

INFER_LEN = 32  # video_depth.py:30-33 ("infer settings, do not change")
OVERLAP = 10
KEYFRAMES = [0, 12, 24, 25, 26, 27, 28, 29, 30, 31]
INTERP_LEN = 8

_MEAN = [0.485, 0.456, 0.406]
_STD = [0.229, 0.224, 0.225]


def effective_input_size(frame_h: int, frame_w: int, input_size: int) -> int:
    """video_depth.py:330-334."""
    ratio = max(frame_h, frame_w) / min(frame_h, frame_w)
    if ratio > 1.78:
        input_size = int(input_size * 1.777 / ratio)
        input_size = round(input_size / 14) * 14
    return input_size


def network_size(frame_h: int, frame_w: int, input_size: int, multiple: int = 14):
    """`Resize.get_size` for keep_aspect_ratio=True, 'lower_bound' (transform.py:51-107).
    Returns (new_h, new_w)."""

    def constrain(x, min_val):
        y = int(np.round(x / multiple) * multiple)
        if y < min_val:
            y = int(np.ceil(x / multiple) * multiple)
        return y

    scale_h = input_size / frame_h
    scale_w = input_size / frame_w
    if scale_w > scale_h:
        scale_h = scale_w
    else:
        scale_w = scale_h
    return constrain(scale_h * frame_h, input_size), constrain(scale_w * frame_w, input_size)


def preprocess_frame(frame_u8: np.ndarray, input_size: int) -> np.ndarray:
    """One frame through Resize(INTER_CUBIC) -> NormalizeImage -> PrepareForNet
    (video_depth.py:336-348,361; transform.py:109-158).  [H,W,3] u8 -> [3,h,w] f32.
    Normalisation happens in float64 (python-list mean/std promote), then casts to f32."""
    import cv2

    img = frame_u8.astype(np.float32) / 255.0
    nh, nw = network_size(img.shape[0], img.shape[1], input_size)
    img = cv2.resize(img, (nw, nh), interpolation=cv2.INTER_CUBIC)
    img = (img - _MEAN) / _STD
    return np.ascontiguousarray(np.transpose(img, (2, 0, 1))).astype(np.float32)


def window_index_map(n_frames: int) -> np.ndarray:
    """Source-frame index of every slot of every window, [n_windows, 32] int64.

    video_depth.py:350-364,374: the list is padded with copies of the last frame;
    window k>0 overwrites its first 10 slots with the previous window's *input*
    slots KEYFRAMES, so the map is pure index arithmetic."""
    step = INFER_LEN - OVERLAP
    append = (step - (n_frames % step)) % step + (INFER_LEN - step)
    padded = list(range(n_frames)) + [n_frames - 1] * append
    rows = []
    prev = None
    for start in range(0, n_frames, step):
        cur = [padded[start + i] for i in range(INFER_LEN)]
        if prev is not None:
            cur[:OVERLAP] = [prev[k] for k in KEYFRAMES]
        rows.append(cur)
        prev = cur
    return np.asarray(rows, dtype=np.int64)


def compute_scale_and_shift(prediction, target, mask):
    """utils/util.py:40-62 (fp32 normal equations, numpy pairwise sums)."""
    prediction = prediction.astype(np.float32)
    target = target.astype(np.float32)
    mask = mask.astype(np.float32)
    a_00 = np.sum(mask * prediction * prediction)
    a_01 = np.sum(mask * prediction)
    a_11 = np.sum(mask)
    b_0 = np.sum(mask * prediction * target)
    b_1 = np.sum(mask * target)
    x_0, x_1 = 1, 0
    det = a_00 * a_11 - a_01 * a_01
    if det != 0:
        x_0 = (a_11 * b_0 - a_01 * b_1) / det
        x_1 = (-a_01 * b_0 + a_00 * b_1) / det
    return x_0, x_1


def blend_frames(pre, post):
    """utils/util.py:65-74: linear cross-fade, weights 0, 1/(n-1), ..., 1 (python floats)."""
    n = len(pre)
    assert n == len(post)
    step = 1.0 / (n - 1)
    w = [0.0] + [i * step for i in range(1, n - 1)] + [1.0]
    return [pre[i] * (1 - w[i]) + post[i] * w[i] for i in range(n)]


def stitch(depth_list, n_frames: int):
    """The alignment / blend loop, video_depth.py:379-417.  `depth_list` is the flat
    list of per-slot depth maps ([H,W] f32), 32 per window.  Returns ([n_frames,H,W],
    [(scale, shift) per window])."""
    depth_list = [d.copy() for d in depth_list]
    aligned, ref_align, st = [], [], []
    align_len = OVERLAP - INTERP_LEN
    kf = KEYFRAMES[:align_len]
    for fid in range(0, len(depth_list), INFER_LEN):
        if len(aligned) == 0:
            aligned += depth_list[:INFER_LEN]
            for k in kf:
                ref_align.append(depth_list[fid + k])
            st.append((1.0, 0.0))
            continue
        cur = [depth_list[fid + i] for i in range(len(kf))]
        scale, shift = compute_scale_and_shift(
            np.concatenate(cur), np.concatenate(ref_align), np.concatenate(np.ones_like(ref_align) == 1)
        )
        st.append((float(scale), float(shift)))
        pre = aligned[-INTERP_LEN:]
        post = depth_list[fid + align_len : fid + OVERLAP]
        for i in range(len(post)):
            post[i] = post[i] * scale + shift
            post[i][post[i] < 0] = 0
        aligned[-INTERP_LEN:] = blend_frames(pre, post)
        for i in range(OVERLAP, INFER_LEN):
            nd = depth_list[fid + i] * scale + shift
            nd[nd < 0] = 0
            aligned.append(nd)
        ref_align = ref_align[:1]
        for k in kf[1:]:
            nd = depth_list[fid + k] * scale + shift
            nd[nd < 0] = 0
            ref_align.append(nd)
    return np.stack(aligned[:n_frames], axis=0), st


def infer_video_depth(frames, target_fps, forward_fn, input_size=518, device="cpu", fp32=True,
                      autocast_dtype=None, return_raw=False):
    """video_depth.py:329-417 with `self.forward` replaced by `forward_fn(x[1,32,3,h,w]) -> [1,32,h,w]`.
    `autocast_dtype`: the reference's dtype-less `torch.autocast(device)` means fp16 on CUDA; pass
    torch.bfloat16 for the bf16 run BASELINE.json asks for."""
    import torch
    import torch.nn.functional as F

    n = frames.shape[0]
    fh, fw = frames[0].shape[:2]
    input_size = effective_input_size(fh, fw, input_size)
    idx = window_index_map(n)
    cache = {}

    def pre(i):
        if i not in cache:
            cache[i] = torch.from_numpy(preprocess_frame(frames[i], input_size))
        return cache[i]

    depth_list = []
    dev_type = torch.device(device).type
    for w in range(idx.shape[0]):
        cur = torch.stack([pre(int(i)) for i in idx[w]], 0).unsqueeze(0).to(device)
        with torch.no_grad():
            if fp32:
                depth = forward_fn(cur)
            else:
                with torch.autocast(device_type=dev_type, dtype=autocast_dtype):
                    depth = forward_fn(cur)
        depth = depth.to(cur.dtype)
        depth = F.interpolate(depth.flatten(0, 1).unsqueeze(1), size=(fh, fw), mode="bilinear", align_corners=True)
        depth_list += [depth[i][0].cpu().numpy() for i in range(depth.shape[0])]
    out, st = stitch(depth_list, n)
    if return_raw:
        return out, target_fps, depth_list, st
    return out, target_fps

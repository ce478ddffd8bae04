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


# --------------------------------------------------------------------------- streaming mode
def stream_schedule(inference_length: int, keyframe_list):
    """The index bookkeeping of `infere_single_image`, video_depth.py:139-185: for each of the first max(keyframes)+1
    predicting steps, the rows of the feature store that form the T-1 context frames (`use_feature_idx`) and the
    positions inside that context used for alignment (`Align_idx`); afterwards the last entry is used for good and the
    store itself is shifted.  Quirks kept: row indices of keyframes that do not exist yet point at still-zero rows."""
    L, nk, max_kf = inference_length, len(keyframe_list), max(keyframe_list)
    dist = [kf + (L - nk) for kf in keyframe_list]
    static = [L - kf if L > kf else i + 1 for i, kf in enumerate(keyframe_list)]
    assert len(static) == len(set(static)), f"Setup leads to duplicates in the keyframes: {static}"
    use, align = [], []
    for fi in range(L - 1, L + max_kf):
        rows = list(range(fi - (L - 1), fi))
        rows[0] = 0  # the first frame is always the reference
        al = [0]
        for i, sk in enumerate(static):
            if sk in rows:
                al.append(rows.index(sk))
            else:
                al.append(i + 1)
                rows[i + 1] = sk if fi - dist[i] <= sk else fi - dist[i]
        use.append(rows)
        align.append(al)
    return use, align


def infere_single_image(frames, target_fps, step_fn, reassemble_fn, input_size=518, device="cpu", fp32=True,
                        inference_length=32, keyframe_list=(0, 12), align_each_new_frame=True, autocast_dtype=None):
    """`VideoDepthAnything.infere_single_image` (warmup=True), video_depth.py:93-327, with the network as parameters:
    reassemble_fn(x[1,1,3,h,w]) -> (layer_1..4) of one frame (encoder + `get_motion_features`);
    step_fn(x, motion_features, pred_depth_idx) -> (depth [1,n,h,w], (layer_1..4 of the new frame))."""
    import torch
    import torch.nn.functional as F

    keyframe_list = list(keyframe_list)
    L, max_kf = inference_length, max(keyframe_list)
    n = frames.shape[0]
    fh, fw = frames[0].shape[:2]
    input_size = effective_input_size(fh, fw, input_size)
    use, align = stream_schedule(L, keyframe_list)
    R = L + max_kf - 1
    move = [i for i in range(R) if i != 1]
    dev_type = torch.device(device).type

    def net(fn, *a):
        with torch.no_grad():
            if fp32:
                return fn(*a)
            with torch.autocast(device_type=dev_type, dtype=autocast_dtype):
                return fn(*a)

    store = None
    depth_list = []
    first = True
    for i in range(n):
        cur = torch.from_numpy(preprocess_frame(frames[i], input_size))[None, None].to(device)
        if i < L - 1:
            feats = net(reassemble_fn, cur)
            if store is None:
                store = [torch.zeros(R, *f.shape[1:]).to(device) for f in feats]
            for s_, f in zip(store, feats):
                s_[i] = f[0]
            continue
        k = min(i - (L - 1), len(use) - 1)
        rows = use[k]
        pred_idx = abs_idx = None
        if align_each_new_frame:
            pred_idx = rows if i == L - 1 else align[k]
            abs_idx = [rows[t] for t in align[k]]
            if i >= L + max_kf:  # the store has been shifted i - (L + max_kf) + 1 times; row 0 never moves
                abs_idx = [0 if a == 0 else a + (i - (L + max_kf)) + 1 for a in abs_idx]
        depth, new = net(step_fn, cur, tuple(s_[rows] for s_ in store), pred_idx)
        for s_, f in zip(store, new):
            if i < R:
                s_[i] = f[0]
            else:
                s_[:-1] = s_[move]
                s_[-1] = f[0]
        depth = depth.to(cur.dtype)
        depth = F.interpolate(depth.flatten(0, 1).unsqueeze(1), size=(fh, fw), mode="bilinear", align_corners=True)
        maps = [depth[j][0].cpu().numpy() for j in range(depth.shape[0])]
        if not align_each_new_frame or first:
            depth_list += maps
            first = False
        else:
            old = [depth_list[j] for j in abs_idx]
            scale, shift = compute_scale_and_shift(np.concatenate(maps[:len(pred_idx)]), np.concatenate(old),
                                                   np.concatenate(np.ones_like(old) == 1))
            depth_list.append(maps[-1] * scale + shift)
    if align_each_new_frame:
        return np.stack(depth_list[1:n], axis=0), target_fps  # the first frame only serves the alignment
    return np.stack(depth_list[:n], axis=0), target_fps


# --------------------------------------------------------------------------- evaluation aligner + metrics
def align_prediction(prediction, ground_truth, valid_depth, max_depth=80.0):
    """utils/align.py:202-236 with its helpers :142-160 restated on plain numpy: the prediction is raw inverse depth, the
    ground truth metric depth; least squares of [p, 1] c = 1/gt (float32 `np.linalg.lstsq`) over valid pixels whose
    ground truth is not zero (inverting the masked array masks those), scale = 1/c0, shift = -c1/c0."""
    prediction = np.asarray(prediction)
    ground_truth = np.asarray(ground_truth)
    valid_depth = np.asarray(valid_depth).astype(bool)
    with np.errstate(divide="ignore"):
        inv = 1 / np.ma.array(ground_truth, mask=~valid_depth)
    mask = ~np.ma.getmaskarray(inv)
    X = prediction[mask][:, None]
    X = np.concatenate([X, np.ones((len(X), 1), dtype=X.dtype)], axis=-1)
    coeffs = np.linalg.lstsq(X, np.ma.getdata(inv)[mask], rcond=None)[0]
    if np.abs(coeffs[0]) <= 0.0:
        scale, shift = float("inf"), 0.0
    else:
        scale, shift = 1 / coeffs[0], -coeffs[1] / coeffs[0]
    a = np.clip((prediction - shift) / scale, 0.0, 1.0)
    a = np.where(a == 0.0, 1e-4, a)
    return np.clip(1.0 / a, 0.0, max_depth), scale, shift


def eval_metrics(prediction, ground_truth, valid_depth):
    """The reference's CSV columns (utils/metrics.py:24-31 over :92-192), numpy."""
    v = np.asarray(valid_depth).astype(bool)
    d, g = np.asarray(prediction)[v], np.asarray(ground_truth)[v]
    with np.errstate(divide="ignore"):
        r = np.maximum(d / g, g / d)
    out = {f"Delta{k}": 1.0 - float((r > 1.25 ** k).mean()) for k in (1, 2, 3)}
    out["SignedRelative"] = float(((d - g) / g).mean())
    out["AbsoluteError"] = float(np.abs(d - g).mean())
    out["AbsoluteRelative"] = float((np.abs(d - g) / g).mean())
    out["MeanSquaredError"] = float(((d - g) ** 2).mean())
    return out

"""ORACLE (test infrastructure only): seeded synthetic inputs shared by goldens and tests."""
import numpy as np
import torch


def synthetic_clip(B: int, T: int, H: int, W: int, seed: int) -> torch.Tensor:
    """N(0,1) network input [B,T,3,H,W] fp32 (normalised-image range)."""
    g = torch.Generator(device="cpu").manual_seed(1000 + seed)
    return torch.randn(B, T, 3, H, W, generator=g)


def synthetic_video(n: int, H: int, W: int, seed: int) -> np.ndarray:
    """uint8 [n,H,W,3] low-pass-filtered noise drifting over time, so that the depth maps are
    smooth but not constant and consecutive windows overlap on similar content."""
    rng = np.random.default_rng(2000 + seed)
    gh, gw = max(2, H // 16 + 2), max(2, W // 16 + 2)
    base = rng.random((n // 8 + 2, gh, gw, 3)).astype(np.float32)
    ys = np.linspace(0, gh - 1.001, H, dtype=np.float32)
    xs = np.linspace(0, gw - 1.001, W, dtype=np.float32)
    y0, x0 = ys.astype(np.int64), xs.astype(np.int64)
    fy, fx = (ys - y0)[:, None, None], (xs - x0)[None, :, None]
    out = np.empty((n, H, W, 3), np.uint8)
    for i in range(n):
        a, t = divmod(i, 8)
        g = base[a] * (1 - t / 8.0) + base[a + 1] * (t / 8.0)
        top = g[y0][:, x0] * (1 - fx) + g[y0][:, x0 + 1] * fx
        bot = g[y0 + 1][:, x0] * (1 - fx) + g[y0 + 1][:, x0 + 1] * fx
        img = top * (1 - fy) + bot * fy
        img = img + rng.normal(0, 0.02, img.shape).astype(np.float32)
        out[i] = np.clip(img * 255.0, 0, 255).astype(np.uint8)
    return out


def synthetic_eval_case(seed: int, n: int = 3, h: int = 40, w: int = 56):
    """Evaluation inputs (eval.py:133-166 shapes): metric ground truth with sensor dropouts (zeros) and an irregular
    validity mask, and a raw inverse-depth prediction that is an affine function of 1/gt plus noise.
    Returns (prediction f32, ground_truth f32, valid bool), each [n, h, w]."""
    import numpy as np

    rng = np.random.default_rng(seed)
    gt = (1.0 + 59.0 * rng.random((n, h, w))).astype(np.float32)
    valid = rng.random((n, h, w)) > 0.2
    gt[rng.random((n, h, w)) < 0.03] = 0.0
    valid &= gt > 0   # the reference's metrics divide by the ground truth of every valid pixel
    pred = (3.0 / np.maximum(gt, 0.5) + 0.2 + 0.02 * rng.standard_normal((n, h, w))).astype(np.float32)
    return pred, gt, valid


def synthetic_depths(seed: int, n: int, h: int, w: int, signed: bool = False) -> np.ndarray:
    """float32 [n, h, w] depth stack for the visualisation tests: smooth ramps plus noise over about two decades;
    `signed` adds negative values (the streaming mode's aligned depths are not clamped, video_depth.py:313)."""
    rng = np.random.default_rng(3000 + seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    d = np.stack([0.05 + 4.0 * ((xx / w + 0.3 * i) % 1.0) * (0.2 + yy / h) for i in range(n)])
    d = d * (1.0 + 0.1 * rng.standard_normal(d.shape)) + (-0.3 if signed else 0.0)
    return (d if signed else np.abs(d)).astype(np.float32)


def synthetic_colour_tables(seed: int):
    """Stand-ins for matplotlib's tables (not in the image): (inferno-like float64 [256, 3], Spectral-like float64 [256, 4]),
    values in [0, 1] including exact 0 and 1 entries."""
    rng = np.random.default_rng(4000 + seed)
    a, b = rng.random((256, 3)), rng.random((256, 4))
    a[0], a[255], b[0], b[255] = 0.0, 1.0, 1.0, 0.0
    return a, b


def synthetic_clip_file(path: str, n: int = 24, h: int = 48, w: int = 64, fps: float = 30.0):
    """Writes the small test clip (moving colour gradients, mp4v) with cv2 -- the same codec stack reads it back."""
    import cv2
    wr = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"mp4v"), fps, (w, h))
    assert wr.isOpened(), path
    yy, xx = np.mgrid[0:h, 0:w]
    for i in range(n):
        wr.write(np.stack([(xx * 4 + i * 5) % 256, (yy * 5 + i * 3) % 256, ((xx + yy) * 2 + i * 7) % 256], -1).astype(np.uint8))
    wr.release()

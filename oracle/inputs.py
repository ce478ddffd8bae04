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

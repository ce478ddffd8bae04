"""Shared helpers for the parity tests (the oracle is imported only here and in the tests)."""
import os

import numpy as np
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel_err_map(d, ref):
    """SURVEY.md section 8c gate: |d - ref| / max(|ref|, 1e-3 * max|ref|), per pixel."""
    d, ref = d.double(), ref.double()
    floor = 1e-3 * ref.abs().max()
    return (d - ref).abs() / torch.clamp(ref.abs(), min=floor)


def abs_rel(d, ref):
    """utils/metrics.py:110-125 semantics with the reference output as ground truth."""
    d, ref = d.double(), ref.double()
    m = ref > 0
    return ((d[m] - ref[m]).abs() / ref[m]).mean().item()


def stats(name, got, ref):
    got, ref = got.double(), ref.double()
    diff = (got - ref).abs()
    scale = ref.abs().max().item() + 1e-30
    s = dict(max_abs=diff.max().item(), max_rel_to_peak=diff.max().item() / scale, mean_abs=diff.mean().item(),
             ref_peak=scale, mismatch_frac=(diff > 0).double().mean().item())
    print(f"[{name}] " + " ".join(f"{k}={v:.3e}" for k, v in s.items()))
    return s


def bf16_close(name, got, ref, ulps=2.0, frac=1e-3):
    """bf16 results: every element within `ulps` bf16 ulps of the (same-rounding-points) reference, except
    a `frac` share that may differ more because an fp32 accumulation-order difference flipped a rounding upstream."""
    g, r = got.double(), ref.double()
    ulp = torch.clamp(r.abs(), min=r.abs().max() * 2 ** -10) * 2 ** -7
    bad = ((g - r).abs() > ulps * ulp).double().mean().item()
    s = stats(name, got, ref)
    assert np.isfinite(s["max_abs"]), name
    assert bad <= frac, f"{name}: {bad:.3e} of elements are more than {ulps} bf16 ulps off"
    return s

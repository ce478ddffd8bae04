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


def bf16_gate(name, ours, ref_bf16, ref_f32, mean_factor=1.5):
    """Whole-network ACCURACY gate of the bf16 build -- not the north-star tolerance, which is printed beside it.

    BASELINE.json states rel-err <= 1e-2 and AbsRel <= 1e-3 against the reference's autocast output.  The reference does
    not meet that against itself: the same reference arithmetic with its GEMMs merely batched differently lands 3e-3 (mean)
    / 3e-2 (max) away (tests/test_gpu_parity.py measures it: a few percent of every bf16 layer's outputs sit close enough
    to a rounding boundary for another fp32 summation order to flip them by one bf16 ulp, and ~20 layers of that add up).
    The strict figures are therefore REPORTED -- "strict gate NOT met" where they are not -- and what is asserted is that
    the bf16 build is as accurate as the reference's own bf16 arithmetic, measured against the reference's fp32 output:
      (a) accuracy:  err(ours, fp32)  <= max(1e-2, 2 x err(ref_bf16, fp32)) at the max (one pixel: a noisy statistic),
                                      <= max(1e-3, mean_factor x err(ref_bf16, fp32)) at the mean -- 1.2 where the network is
                                      not chaotic (trained-like weights); 1.5 with the random recipe, whose 12 / 24 blocks
                                      amplify each run's per-block rounding error (equal to the reference's, stage by stage:
                                      tests/test_gpu_parity.py) by a factor that varies from input to input
      (b) agreement: err(ours, ref_bf16) <= max(1e-2, 2 x err(ref_bf16, fp32)) at the max; AbsRel <= max(1e-3, 2 x) -- two
          independent bf16 evaluations differ by about sqrt(2) x the deviation of either from the fp32 result
    Rounding-point fidelity itself is pinned stage by stage with teacher-forced inputs (tests/test_gpu_parity.py) and kernel
    by kernel (test_gpu_kernels.py / test_gpu_tc.py, <= 2 bf16 ulps)."""
    e_ref, e_ours, e_x = rel_err_map(ref_bf16, ref_f32), rel_err_map(ours, ref_f32), rel_err_map(ours, ref_bf16)
    ar_ref, ar_x = abs_rel(ref_bf16, ref_f32), abs_rel(ours, ref_bf16)
    r = dict(ref_max=e_ref.max().item(), ref_mean=e_ref.mean().item(), ours_max=e_ours.max().item(),
             ours_mean=e_ours.mean().item(), x_max=e_x.max().item(), x_mean=e_x.mean().item(), absrel_ref=ar_ref, absrel_x=ar_x)
    strict = r["x_max"] <= 1e-2 and ar_x <= 1e-3
    print(f"[{name}] ours-vs-ref_bf16 max {r['x_max']:.3e} mean {r['x_mean']:.3e} AbsRel {ar_x:.3e} | ref_bf16-vs-fp32 max "
          f"{r['ref_max']:.3e} mean {r['ref_mean']:.3e} AbsRel {ar_ref:.3e} | ours-vs-fp32 max {r['ours_max']:.3e} mean "
          f"{r['ours_mean']:.3e} | strict 1e-2/1e-3 gate {'met' if strict else 'NOT met (nor by the reference against itself)'}")
    assert np.isfinite(r["x_max"]) and np.isfinite(r["ours_max"]), name
    assert r["ours_max"] <= max(1e-2, 2.0 * r["ref_max"]), (name, r)
    assert r["ours_mean"] <= max(1e-3, mean_factor * r["ref_mean"]), (name, r)
    assert r["x_max"] <= max(1e-2, 2.0 * r["ref_max"]), (name, r)
    assert ar_x <= max(1e-3, 2.0 * ar_ref), (name, r)
    return r

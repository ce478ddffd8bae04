"""Evaluation aligner and metrics on the device (reference utils/align.py:151-236, utils/metrics.py:92-192, as used by
eval.py:163-181): the masked least-squares fit of the raw (inverse-depth) prediction to 1/ground-truth, the aligned
metric depth, and the seven error statistics of the reference's CSV row, in two passes over device-resident arrays.

`align_prediction` keeps the reference's name, arguments and return value (aligned prediction, scale, shift); arrays may
be numpy arrays, CPU tensors or CUDA tensors, and the aligned prediction comes back in the kind that went in.
"""
from __future__ import annotations

import numpy as np
import torch

from . import lib as L

METRIC_NAMES = ("Delta1", "Delta2", "Delta3", "SignedRelative", "AbsoluteError", "AbsoluteRelative", "MeanSquaredError")


def _dev(x, dtype, device):
    t = torch.from_numpy(np.ascontiguousarray(x)) if isinstance(x, np.ndarray) else x
    return t.to(device=device, dtype=dtype, non_blocking=True).contiguous()


class DeviceEvaluator:
    def __init__(self, device="cuda"):
        self.device = torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        L.require_device(self.device)
        self.lib = L.load()
        self.scratch = torch.zeros(5, dtype=torch.float64, device=self.device)
        self.sums = torch.zeros(8, dtype=torch.float64, device=self.device)
        self.coef = torch.zeros(2, dtype=torch.float64, device=self.device)

    def fit(self, prediction, ground_truth, valid_depth):
        """(scale, shift) of utils/align.py:151-160 as a device tensor [2]; inputs already on the device."""
        with torch.cuda.device(self.device):
            L.check(self.lib.vda_eval_fit(prediction.data_ptr(), ground_truth.data_ptr(), valid_depth.data_ptr(),
                                          prediction.numel(), self.scratch.data_ptr(), self.coef.data_ptr(), 3, L.stream_ptr()),
                    "vda_eval_fit")
        return self.coef

    def apply(self, prediction, ground_truth, valid_depth, coef, max_depth, want_aligned=True):
        """Aligned metric depth (or None) and the metric sums [8] (device), utils/align.py:228-236 + utils/metrics.py."""
        aligned = torch.empty_like(prediction) if want_aligned else None
        with torch.cuda.device(self.device):
            L.check(self.lib.vda_eval_apply(prediction.data_ptr(), ground_truth.data_ptr(), valid_depth.data_ptr(),
                                            coef.data_ptr(), float(max_depth), L._ptr(aligned), prediction.numel(),
                                            self.sums.data_ptr(), L.stream_ptr()), "vda_eval_apply")
        return aligned, self.sums

    @staticmethod
    def metrics_from_sums(sums):
        """The reference's CSV columns (utils/metrics.py:9,24-31) from the eight sums."""
        s = [float(v) for v in sums.tolist()]
        n = s[0]
        if n == 0:
            return {k: float("nan") for k in METRIC_NAMES}
        return {"Delta1": 1.0 - s[5] / n, "Delta2": 1.0 - s[6] / n, "Delta3": 1.0 - s[7] / n, "SignedRelative": s[3] / n,
                "AbsoluteError": s[1] / n, "AbsoluteRelative": s[2] / n, "MeanSquaredError": s[4] / n}


_evaluators = {}


def _evaluator(device):
    key = str(torch.device(device))
    if key not in _evaluators:
        _evaluators[key] = DeviceEvaluator(device)
    return _evaluators[key]


def align_prediction(prediction, ground_truth, valid_depth, max_depth=80.0, device="cuda", return_metrics=False):
    """Drop-in for utils/align.py:202-236.  Returns (aligned prediction, scale, shift) -- plus the metric dict of
    utils/metrics.py's CSV row when `return_metrics` -- computed on `device`."""
    ev = _evaluator(prediction.device if isinstance(prediction, torch.Tensor) and prediction.is_cuda else device)
    as_numpy = isinstance(prediction, np.ndarray)
    on_cpu = as_numpy or not prediction.is_cuda
    p = _dev(prediction, torch.float32, ev.device)
    g = _dev(ground_truth, torch.float32, ev.device)
    v = _dev(valid_depth, torch.bool, ev.device).view(torch.uint8)
    coef = ev.fit(p, g, v)
    aligned, sums = ev.apply(p, g, v, coef, max_depth)
    scale, shift = (np.float64(x) for x in coef.tolist())
    if on_cpu:
        aligned = aligned.cpu()
        if as_numpy:
            aligned = aligned.numpy()
    if return_metrics:
        return aligned, scale, shift, ev.metrics_from_sums(sums)
    return aligned, scale, shift

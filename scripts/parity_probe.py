#!/usr/bin/env python
"""Where does the bf16 build's deviation from the fp32 reference come from?  Whole-network figures for variants of the plan
(exact-rounding SIMT attention instead of the tcgen05 flash attention, TF32 output head instead of fp16 operands) on one
configuration.  Test infrastructure (uses the oracle)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.inputs import synthetic_clip  # noqa: E402
from tests.test_gpu_parity import get_model, oracle_bf16, oracle_fp32, release  # noqa: E402
from tests.util import rel_err_map  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--enc", default="vitl")
    ap.add_argument("--T", type=int, default=32)
    ap.add_argument("--H", type=int, default=518)
    ap.add_argument("--W", type=int, default=518)
    ap.add_argument("--style", default="random")
    a = ap.parse_args()
    m, sd = get_model(a.enc, a.style)
    x = synthetic_clip(1, a.T, a.H, a.W, 33).to("cuda")
    ref, f32 = oracle_bf16(sd, x, a.enc), oracle_fp32(sd, x, a.enc)
    print(f"{a.enc} T{a.T} {a.H}x{a.W} {a.style}: reference bf16 vs fp32 mean {rel_err_map(ref, f32).mean().item():.3e}")
    for name, attn_simt, tf32 in (("default", False, False), ("simt attention", True, False), ("tf32 head", False, True),
                                   ("simt attention + tf32 head", True, True)):
        os.environ["VDA_HEAD_TF32"] = "1" if tf32 else "0"
        m.attn_simt = attn_simt
        release(m)
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            d = m(x)
        print(f"  {name:28s} ours vs fp32 mean {rel_err_map(d, f32).mean().item():.3e} max {rel_err_map(d, f32).max().item():.3e} | "
              f"ours vs reference bf16 mean {rel_err_map(d, ref).mean().item():.3e}")
    m.attn_simt = False


if __name__ == "__main__":
    main()

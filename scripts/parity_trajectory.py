#!/usr/bin/env python
"""Free-running trajectories: how far the bf16 build and the reference's bf16 autocast run are from the fp32 reference at
every block boundary of the encoder and at the head's named maps.  Test infrastructure (uses the oracle)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import model as omodel  # noqa: E402
from oracle.inputs import synthetic_clip  # noqa: E402
from oracle.weights import ENCODER_DIMS  # noqa: E402
from tests.test_gpu_parity import exact_fp32, get_model, read_dense, read_padded, stage_end  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--enc", default="vitl")
    ap.add_argument("--T", type=int, default=8)
    ap.add_argument("--H", type=int, default=518)
    ap.add_argument("--W", type=int, default=518)
    ap.add_argument("--style", default="random")
    a = ap.parse_args()
    enc, T, H, W = a.enc, a.T, a.H, a.W
    m, sd = get_model(enc, a.style)
    x = synthetic_clip(1, T, H, W, 33).to("cuda")
    cb, cf = {"want_blocks": True}, {"want_blocks": True}
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        rb = omodel.forward(sd, x, enc, collect=cb)
    with exact_fp32(), torch.no_grad():
        rf = omodel.forward(sd, x, enc, collect=cf)
    plan = m.plan(1, T, H, W, torch.device("cuda", torch.cuda.current_device()), True)
    plan.x_in.copy_(x.flatten(0, 1))
    dims = ENCODER_DIMS[enc]
    D, depth = dims["dim"], dims["depth"]
    ph, pw = H // 14, W // 14
    ntok = ph * pw + 1

    class Err(float):
        bias = 0.0

        def __format__(self, spec):
            return f"{float(self):{spec}} (signed {self.bias:+.2e}, along-signal {self.along:+.2e})"

    def err(a_, f):
        a_, f = a_.float(), f.float()
        e = Err((a_ - f).abs().sum().item() / f.abs().sum().item())
        e.bias = (a_ - f).sum().item() / f.abs().sum().item()                       # mean signed deviation
        e.along = ((a_ - f) * torch.sign(f)).sum().item() / f.abs().sum().item()    # > 0: magnitudes grow, < 0: they shrink
        return e

    plan.run_range(0, plan.marks["blk0"])
    print(f"{'tokens0':28s} ours {err(plan.bufs['resid'].view(T, ntok, D), cf['blk0_in']):.3e}  reference bf16 {err(cb['blk0_in'], cf['blk0_in']):.3e}")
    for i in range(depth):
        s = plan.marks[f"blk{i}"]
        e = plan.marks[f"blk{i + 1}"] if i + 1 < depth else plan.marks["enc_end"]
        plan.run_range(s, e)
        print(f"{'after block ' + str(i):28s} ours {err(plan.bufs['resid'].view(T, ntok, D), cf[f'blk{i + 1}_in']):.3e}  "
              f"reference bf16 {err(cb[f'blk{i + 1}_in'], cf[f'blk{i + 1}_in']):.3e}")
    plan.run_range(plan.marks["enc_end"], len(plan.ops))
    B_ = plan.bufs
    h1, w1, h2, w2, h3, w3, h4, w4 = 4 * ph, 4 * pw, 2 * ph, 2 * pw, ph, pw, (ph + 1) // 2, (pw + 1) // 2
    maps = [("layer_1", read_padded(B_["layer_1"], T, h1, w1)), ("layer_3", read_dense(B_["layer_3"], T, h3, w3)),
            ("mm0", read_padded(B_["mm0"], T, h3, w3)), ("mm1", read_padded(B_["mm1"], T, h4, w4)),
            ("path_4", read_padded(B_["path_4"], T, h3, w3)), ("path_3", read_padded(B_["path_3"], T, h2, w2)),
            ("path_2", read_padded(B_["path_2"], T, h1, w1)), ("path_1", read_padded(B_["path_1"], T, 2 * h1, 2 * w1))]
    for name, ours in maps:
        print(f"{name:28s} ours {err(ours, cf[name]):.3e}  reference bf16 {err(cb[name], cf[name]):.3e}")
    print(f"{'depth':28s} ours {err(plan.depth.view(1, T, H, W), rf):.3e}  reference bf16 {err(rb, rf):.3e}")


if __name__ == "__main__":
    main()

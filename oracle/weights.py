"""ORACLE (test infrastructure only -- never imported by the product path).

Parameter inventory and the fixed-seed weight recipe shared by the oracle, the
golden-vector generator and the GPU parity tests.

The reference ships no checkpoints (`/root/reference/.gitignore:3`,
`get_weights.sh:3-5`), so parity runs on seeded random weights.  The reference's
default initialisation is degenerate for testing:
`proj_out` of every motion module is zero (`motion_module/motion_module.py:61-62`),
LayerScale gammas are all 1.0 (`dinov2.py:409`), the cls token has std 1e-6
(`dinov2.py:174`).  The recipe below therefore draws *every* tensor from a
non-trivial distribution, keyed by tensor name so the result does not depend on
construction order, torch module code, or the reference being importable.

`param_spec()` restates the reference's state-dict layout
(`video_depth.py:36-56`, `dinov2.py:339-378`, `dpt.py:47-124`,
`dpt_temporal.py:35-51`, `motion_module.py:74-107,136-171,189-207`); it is
checked key-for-key against the real reference by `scripts/make_goldens.py`
(run in the build container) and against `tests/golden/state_dict_keys_*.json`
by the CPU test-suite.
"""
from __future__ import annotations

import math
import zlib
from collections import OrderedDict

import torch

MODEL_CONFIGS = {
    # video_depth.py:48-51 / run.py:74-77
    "vits": dict(encoder="vits", features=64, out_channels=[48, 96, 192, 384]),
    "vitl": dict(encoder="vitl", features=256, out_channels=[256, 512, 1024, 1024]),
}

ENCODER_DIMS = {
    # dinov2.py:339-378 (embed_dim, depth, heads); taps video_depth.py:48-51
    "vits": dict(dim=384, depth=12, heads=6, taps=[2, 5, 8, 11]),
    "vitl": dict(dim=1024, depth=24, heads=16, taps=[4, 11, 17, 23]),
}


def param_spec(encoder: str, num_frames: int = 32, pe: str = "ape") -> "OrderedDict[str, tuple]":
    """name -> shape for every state-dict entry, in the reference's own order.  pe='rope' models have no
    `pos_encoder.pe` buffers: freqs_cis is a plain attribute (motion_module.py:226-240)."""
    cfg = MODEL_CONFIGS[encoder]
    enc = ENCODER_DIMS[encoder]
    D, depth = enc["dim"], enc["depth"]
    F = cfg["features"]
    oc = cfg["out_channels"]
    s: "OrderedDict[str, tuple]" = OrderedDict()
    p = "pretrained."
    s[p + "cls_token"] = (1, 1, D)
    s[p + "pos_embed"] = (1, 37 * 37 + 1, D)
    s[p + "mask_token"] = (1, D)
    s[p + "patch_embed.proj.weight"] = (D, 3, 14, 14)
    s[p + "patch_embed.proj.bias"] = (D,)
    for i in range(depth):
        b = f"{p}blocks.{i}."
        s[b + "norm1.weight"] = (D,)
        s[b + "norm1.bias"] = (D,)
        s[b + "attn.qkv.weight"] = (3 * D, D)
        s[b + "attn.qkv.bias"] = (3 * D,)
        s[b + "attn.proj.weight"] = (D, D)
        s[b + "attn.proj.bias"] = (D,)
        s[b + "ls1.gamma"] = (D,)
        s[b + "norm2.weight"] = (D,)
        s[b + "norm2.bias"] = (D,)
        s[b + "mlp.fc1.weight"] = (4 * D, D)
        s[b + "mlp.fc1.bias"] = (4 * D,)
        s[b + "mlp.fc2.weight"] = (D, 4 * D)
        s[b + "mlp.fc2.bias"] = (D,)
        s[b + "ls2.gamma"] = (D,)
    s[p + "norm.weight"] = (D,)
    s[p + "norm.bias"] = (D,)

    h = "head."
    for i in range(4):
        s[f"{h}projects.{i}.weight"] = (oc[i], D, 1, 1)
        s[f"{h}projects.{i}.bias"] = (oc[i],)
    s[h + "resize_layers.0.weight"] = (oc[0], oc[0], 4, 4)
    s[h + "resize_layers.0.bias"] = (oc[0],)
    s[h + "resize_layers.1.weight"] = (oc[1], oc[1], 2, 2)
    s[h + "resize_layers.1.bias"] = (oc[1],)
    s[h + "resize_layers.3.weight"] = (oc[3], oc[3], 3, 3)
    s[h + "resize_layers.3.bias"] = (oc[3],)
    sc = h + "scratch."
    for i in range(4):
        s[f"{sc}layer{i + 1}_rn.weight"] = (F, oc[i], 3, 3)
    for i in range(1, 5):
        r = f"{sc}refinenet{i}."
        s[r + "out_conv.weight"] = (F, F, 1, 1)
        s[r + "out_conv.bias"] = (F,)
        for u in (1, 2):
            for c in (1, 2):
                s[f"{r}resConfUnit{u}.conv{c}.weight"] = (F, F, 3, 3)
                s[f"{r}resConfUnit{u}.conv{c}.bias"] = (F,)
    s[sc + "output_conv1.weight"] = (F // 2, F, 3, 3)
    s[sc + "output_conv1.bias"] = (F // 2,)
    s[sc + "output_conv2.0.weight"] = (32, F // 2, 3, 3)
    s[sc + "output_conv2.0.bias"] = (32,)
    s[sc + "output_conv2.2.weight"] = (1, 32, 1, 1)
    s[sc + "output_conv2.2.bias"] = (1,)
    for m, C in enumerate([oc[2], oc[3], F, F]):
        t = f"{h}motion_modules.{m}.temporal_transformer."
        s[t + "norm.weight"] = (C,)
        s[t + "norm.bias"] = (C,)
        s[t + "proj_in.weight"] = (C, C)
        s[t + "proj_in.bias"] = (C,)
        tb = t + "transformer_blocks.0."
        for a in (0, 1):
            ab = f"{tb}attention_blocks.{a}."
            s[ab + "to_q.weight"] = (C, C)
            s[ab + "to_k.weight"] = (C, C)
            s[ab + "to_v.weight"] = (C, C)
            s[ab + "to_out.0.weight"] = (C, C)
            s[ab + "to_out.0.bias"] = (C,)
            if pe == "ape":
                s[ab + "pos_encoder.pe"] = (1, num_frames, C)
        for a in (0, 1):
            s[f"{tb}norms.{a}.weight"] = (C,)
            s[f"{tb}norms.{a}.bias"] = (C,)
        s[tb + "ff.net.0.proj.weight"] = (8 * C, C)
        s[tb + "ff.net.0.proj.bias"] = (8 * C,)
        s[tb + "ff.net.2.weight"] = (C, 4 * C)
        s[tb + "ff.net.2.bias"] = (C,)
        s[tb + "ff_norm.weight"] = (C,)
        s[tb + "ff_norm.bias"] = (C,)
        s[t + "proj_out.weight"] = (C, C)
        s[t + "proj_out.bias"] = (C,)
    return s


def sinusoidal_pe(max_len: int, d_model: int) -> torch.Tensor:
    """`PositionalEncoding.__init__` buffer, motion_module.py:197-203."""
    position = torch.arange(max_len).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, d_model, 2) * (-math.log(10000.0) / d_model))
    pe = torch.zeros(1, max_len, d_model)
    pe[0, :, 0::2] = torch.sin(position * div_term)
    pe[0, :, 1::2] = torch.cos(position * div_term)
    return pe


def _gen(name: str, seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(name.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    return g


def make_state_dict(encoder: str, seed: int = 0, num_frames: int = 32, pe: str = "ape",
                    style: str = "random") -> "OrderedDict[str, torch.Tensor]":
    """The fixed-seed weight recipe, fp32 CPU tensors.

    Weights ~ N(0, g/sqrt(fan_in)); biases ~ N(0, 0.05); every norm gets a
    non-trivial affine; LayerScale gamma ~ U(0.5, 1.5); proj_out is *not* zero;
    the last 1x1 conv is mostly positive so the depth map is not half zeros.

    style="random" (the goldens): every residual branch is as large as the stream it is added to, so a 12 / 24-block
    encoder amplifies one flipped bf16 rounding into an independent realisation of the network.
    style="trained_like": the statistics of a trained checkpoint instead -- LayerScale gamma ~ 0.1 x U(0.5, 1.5) (DINOv2
    initialises it at 1e-5 and trains it up to a fraction of one), softmax logits without the extra gain, motion-module
    `proj_out` (zero-initialised in the reference, motion_module.py:61-62) at a quarter of the random scale: each branch
    is a perturbation of its residual stream, as in a trained network, and rounding noise averages instead of
    amplifying.  Everything else is identical; every branch of the path stays live.
    """
    assert style in ("random", "trained_like")
    smooth = style == "trained_like"
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for name, shape in param_spec(encoder, num_frames, pe).items():
        g = _gen(name, seed)
        leaf = name.rsplit(".", 1)[-1]
        if leaf == "pe":
            t = sinusoidal_pe(shape[1], shape[2])
        elif leaf == "gamma":
            t = torch.rand(shape, generator=g) + 0.5
            if smooth:
                t = t * 0.1
        elif leaf in ("cls_token", "pos_embed"):
            t = torch.randn(shape, generator=g) * 0.02
        elif leaf == "mask_token":
            t = torch.zeros(shape)
        elif len(shape) == 1 and leaf == "weight":  # LayerNorm / GroupNorm scale
            t = 0.8 + 0.4 * torch.rand(shape, generator=g)
        elif leaf == "bias":
            t = torch.randn(shape, generator=g) * 0.05
            if name.endswith("output_conv2.2.bias"):
                t = torch.full(shape, 0.05)
        else:
            assert leaf == "weight" and len(shape) in (2, 4), name
            if "resize_layers.0" in name or "resize_layers.1" in name:
                fan_in = shape[0]  # ConvTranspose2d: [Cin, Cout, k, k], each output sees Cin
            else:
                fan_in = 1
                for d in shape[1:]:
                    fan_in *= d
            gain = 1.0
            if name.endswith("attn.qkv.weight") or name.endswith("to_q.weight") or name.endswith("to_k.weight"):
                gain = 1.0 if smooth else 1.3  # peakier softmax than the trunc_normal(0.02) default
            if "resConfUnit" in name or "output_conv" in name or "_rn." in name:
                gain = 1.4  # ReLU-fed convs: keep the signal alive through the fusion stack
            if smooth and ".proj_out." in name:
                gain = 0.25
            t = torch.randn(shape, generator=g) * (gain / math.sqrt(fan_in))
            if name.endswith("output_conv2.2.weight"):
                t = t.abs() * 0.8 + t * 0.2
        sd[name] = t.contiguous()
    return sd

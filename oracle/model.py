"""ORACLE (test infrastructure only -- never imported by the product path).

A functional PyTorch restatement of the reference's inference forward
(`VideoDepthAnything.forward`, /root/reference/video_depth_anything/video_depth.py:58-65)
operating directly on a state-dict.  It uses the same torch ops, in the same
order, as the reference modules, so that

  * on CPU in fp32 it reproduces the reference bit-for-bit up to op-order
    (pinned by `scripts/make_goldens.py` against the real reference, and by
    `tests/test_oracle_cpu.py` against the committed goldens), and
  * on a B200 under `torch.autocast('cuda', dtype=torch.bfloat16)` it takes the
    same autocast cast decisions as the reference (autocast acts per dispatcher
    op, not per nn.Module), which is what the bf16 parity gate compares with.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py --impl reference` /
`cpu_baseline` may import this file.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from .weights import ENCODER_DIMS, MODEL_CONFIGS

PATCH = 14


# --------------------------------------------------------------------------- encoder
def interpolate_pos_embed(pos_embed: torch.Tensor, npatch: int, H: int, W: int, out_dtype) -> torch.Tensor:
    """dinov2.py:179-210.  NB the reference calls this with (w, h) = (H, W) of the
    image (dinov2.py:213 names them backwards); the identity shortcut needs H == W."""
    N = pos_embed.shape[1] - 1
    if npatch == N and H == W:
        return pos_embed
    pe = pos_embed.float()
    cls_pe, patch_pe = pe[:, 0], pe[:, 1:]
    dim = pe.shape[-1]
    h0, w0 = H // PATCH + 0.1, W // PATCH + 0.1  # interpolate_offset = 0.1 (dinov2.py:412)
    s = math.sqrt(N)
    patch_pe = F.interpolate(
        patch_pe.reshape(1, int(s), int(s), dim).permute(0, 3, 1, 2),
        scale_factor=(float(h0) / s, float(w0) / s),
        mode="bicubic",
        antialias=False,
    )
    assert patch_pe.shape[-2] == int(h0) and patch_pe.shape[-1] == int(w0)
    patch_pe = patch_pe.permute(0, 2, 3, 1).reshape(1, -1, dim)
    return torch.cat((cls_pe.unsqueeze(0), patch_pe), dim=1).to(out_dtype)


def encoder_block(sd, p: str, x: torch.Tensor, heads: int) -> torch.Tensor:
    """block.py:83-87,105-106 with attention.py:49-62 (no-xformers fallback) and mlp.py:35-41."""
    B, N, C = x.shape
    y = F.layer_norm(x, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"], 1e-6)
    qkv = F.linear(y, sd[p + "attn.qkv.weight"], sd[p + "attn.qkv.bias"])
    qkv = qkv.reshape(B, N, 3, heads, C // heads).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0] * (C // heads) ** -0.5, qkv[1], qkv[2]
    a = (q @ k.transpose(-2, -1)).softmax(dim=-1)
    y = (a @ v).transpose(1, 2).reshape(B, N, C)
    y = F.linear(y, sd[p + "attn.proj.weight"], sd[p + "attn.proj.bias"])
    x = x + y * sd[p + "ls1.gamma"]
    y = F.layer_norm(x, (C,), sd[p + "norm2.weight"], sd[p + "norm2.bias"], 1e-6)
    y = F.linear(y, sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"])
    y = F.gelu(y)
    y = F.linear(y, sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"])
    return x + y * sd[p + "ls2.gamma"]


def encoder_taps(sd, x: torch.Tensor, encoder: str, collect=None):
    """`get_intermediate_layers(x, taps, return_class_token=True)`, dinov2.py:297-321,
    271-281, 212-231.  x: [BT,3,H,W].  Returns 4 normalised patch-token tensors [BT, N-1, D]."""
    enc = ENCODER_DIMS[encoder]
    BT, _, H, W = x.shape
    assert H % PATCH == 0 and W % PATCH == 0  # patch_embed.py:73-74
    p = "pretrained."
    t = F.conv2d(x, sd[p + "patch_embed.proj.weight"], sd[p + "patch_embed.proj.bias"], stride=PATCH)
    t = t.flatten(2).transpose(1, 2)
    t = torch.cat((sd[p + "cls_token"].expand(BT, -1, -1), t), dim=1)
    t = t + interpolate_pos_embed(sd[p + "pos_embed"], t.shape[1] - 1, H, W, t.dtype)
    if collect is not None:
        collect["tokens0"] = t
    taps = []
    for i in range(enc["depth"]):
        if collect is not None and collect.get("want_blocks"):
            collect[f"blk{i}_in"] = t
        t = encoder_block(sd, f"{p}blocks.{i}.", t, enc["heads"])
        if i in enc["taps"]:
            taps.append(t)
    if collect is not None and collect.get("want_blocks"):
        collect[f"blk{enc['depth']}_in"] = t
    D = t.shape[-1]
    taps = [F.layer_norm(o, (D,), sd[p + "norm.weight"], sd[p + "norm.bias"], 1e-6)[:, 1:] for o in taps]
    if collect is not None:
        for i, o in enumerate(taps):
            collect[f"tap{i}"] = o
    return taps


# --------------------------------------------------------------------------- motion module
def rope_freqs_cis(dim: int, end: int, theta: float = 10000.0) -> torch.Tensor:
    """`precompute_freqs_cis`, motion_module/attention.py:403-408 (complex64 [end, dim/2])."""
    freqs = 1.0 / (theta ** (torch.arange(0, dim, 2)[: (dim // 2)].float() / dim))
    t = torch.arange(end, dtype=torch.float32)
    return torch.polar(torch.ones(end, dim // 2), torch.outer(t, freqs))


def rope_rotate(x: torch.Tensor, freqs_cis: torch.Tensor) -> torch.Tensor:
    """`apply_rotary_emb` for one operand, motion_module/attention.py:419-429: x [n, T, c] viewed as c/2 complex
    numbers per (n, t), multiplied by freqs_cis[t] in fp32, cast back to x's dtype."""
    xc = torch.view_as_complex(x.float().reshape(*x.shape[:-1], -1, 2).contiguous())
    return torch.view_as_real(xc * freqs_cis.view(1, xc.shape[1], xc.shape[2])).flatten(2).type_as(x)


def temporal_attention(sd, p: str, h: torch.Tensor, T: int, heads: int = 8) -> torch.Tensor:
    """motion_module.py:247-335 + attention.py:93-111,182-211.  h: [(b f), d, c].  The positional scheme follows the
    state dict: a `pos_encoder.pe` buffer means pe='ape' (added to the input, :255-256); its absence means pe='rope'
    (query and key rotated after their projections, :290-293)."""
    BF, d, c = h.shape
    b = BF // T
    h = h.reshape(b, T, d, c).permute(0, 2, 1, 3).reshape(b * d, T, c)  # "(b f) d c -> (b d) f c"
    ape = (p + "pos_encoder.pe") in sd
    if ape:
        pe = sd[p + "pos_encoder.pe"]
        h = h + pe[:, :T].to(h.dtype)
    q = F.linear(h, sd[p + "to_q.weight"])
    k = F.linear(h, sd[p + "to_k.weight"])
    v = F.linear(h, sd[p + "to_v.weight"])
    if not ape:
        fc = rope_freqs_cis(c, T).to(q.device)  # == precompute_freqs_cis(c, temporal_max_len)[:T]
        q, k = rope_rotate(q, fc), rope_rotate(k, fc)
    dh = c // heads

    def split(t):
        return t.reshape(b * d, T, heads, dh).permute(0, 2, 1, 3).reshape(b * d * heads, T, dh)

    q, k, v = split(q), split(k), split(v)
    s = torch.baddbmm(
        torch.empty(q.shape[0], T, T, dtype=q.dtype, device=q.device), q, k.transpose(-1, -2), beta=0, alpha=dh**-0.5
    )
    a = s.softmax(dim=-1).to(v.dtype)
    o = torch.bmm(a, v)
    o = o.reshape(b * d, heads, T, dh).permute(0, 2, 1, 3).reshape(b * d, T, c)
    o = F.linear(o, sd[p + "to_out.0.weight"], sd[p + "to_out.0.bias"])
    return o.reshape(b, d, T, c).permute(0, 2, 1, 3).reshape(BF, d, c)  # "(b d) f c -> (b f) d c"


def temporal_module(sd, p: str, x: torch.Tensor, T: int) -> torch.Tensor:
    """`TemporalModule.forward` on frames-flattened NCHW input [(b f), C, h, w]
    (motion_module.py:64-69,108-133,173-186; FeedForward/GEGLU attention.py:335-338,382-384)."""
    BF, C, hh, ww = x.shape
    p = p + "temporal_transformer."
    res = x
    y = F.group_norm(x, 32, sd[p + "norm.weight"], sd[p + "norm.bias"], 1e-6)
    y = y.permute(0, 2, 3, 1).reshape(BF, hh * ww, C)
    y = F.linear(y, sd[p + "proj_in.weight"], sd[p + "proj_in.bias"])
    tb = p + "transformer_blocks.0."
    for a in (0, 1):
        n = F.layer_norm(y, (C,), sd[f"{tb}norms.{a}.weight"], sd[f"{tb}norms.{a}.bias"], 1e-5)
        y = temporal_attention(sd, f"{tb}attention_blocks.{a}.", n, T) + y
    n = F.layer_norm(y, (C,), sd[tb + "ff_norm.weight"], sd[tb + "ff_norm.bias"], 1e-5)
    val, gate = F.linear(n, sd[tb + "ff.net.0.proj.weight"], sd[tb + "ff.net.0.proj.bias"]).chunk(2, dim=-1)
    n = F.linear(val * F.gelu(gate), sd[tb + "ff.net.2.weight"], sd[tb + "ff.net.2.bias"])
    y = n + y
    y = F.linear(y, sd[p + "proj_out.weight"], sd[p + "proj_out.bias"])
    y = y.reshape(BF, hh, ww, C).permute(0, 3, 1, 2)
    return y + res


# --------------------------------------------------------------------------- DPT head
def _rcu(sd, p: str, x: torch.Tensor) -> torch.Tensor:
    """ResidualConvUnit, blocks.py:68-91 (non-inplace ReLU: the skip is the raw input)."""
    y = F.conv2d(F.relu(x), sd[p + "conv1.weight"], sd[p + "conv1.bias"], padding=1)
    y = F.conv2d(F.relu(y), sd[p + "conv2.weight"], sd[p + "conv2.bias"], padding=1)
    return y + x


def _fusion(sd, p: str, x0: torch.Tensor, x1, size) -> torch.Tensor:
    """FeatureFusionBlock.forward, blocks.py:135-162."""
    out = x0
    if x1 is not None:
        out = out + _rcu(sd, p + "resConfUnit1.", x1)
    out = _rcu(sd, p + "resConfUnit2.", out)
    if size is None:
        out = F.interpolate(out, scale_factor=2, mode="bilinear", align_corners=True)
    else:
        out = F.interpolate(out, size=size, mode="bilinear", align_corners=True)
    return F.conv2d(out, sd[p + "out_conv.weight"], sd[p + "out_conv.bias"])


def reassemble(sd, taps, ph: int, pw: int):
    """projects + resize_layers: `DPTHeadTemporal.get_motion_features`, dpt_temporal.py:101-130 (== :55-69)."""
    h = "head."
    layers = []
    for i, x in enumerate(taps):
        BT, _, D = x.shape
        x = x.permute(0, 2, 1).reshape(BT, D, ph, pw)
        x = F.conv2d(x, sd[f"{h}projects.{i}.weight"], sd[f"{h}projects.{i}.bias"])
        if i == 0:
            x = F.conv_transpose2d(x, sd[h + "resize_layers.0.weight"], sd[h + "resize_layers.0.bias"], stride=4)
        elif i == 1:
            x = F.conv_transpose2d(x, sd[h + "resize_layers.1.weight"], sd[h + "resize_layers.1.bias"], stride=2)
        elif i == 3:
            x = F.conv2d(x, sd[h + "resize_layers.3.weight"], sd[h + "resize_layers.3.bias"], stride=2, padding=1)
        layers.append(x)
    return layers


def _output_convs(sd, p1: torch.Tensor, ph: int, pw: int) -> torch.Tensor:
    """output_conv1 -> bilinear -> fp32 island output_conv2 (dpt_temporal.py:91-99 == :230-240,249-258)."""
    sc = "head.scratch."
    out = F.conv2d(p1, sd[sc + "output_conv1.weight"], sd[sc + "output_conv1.bias"], padding=1)
    out = F.interpolate(out, (ph * PATCH, pw * PATCH), mode="bilinear", align_corners=True)
    ori = out.dtype
    with torch.autocast(device_type="cuda", enabled=False):
        o = F.conv2d(out.float(), sd[sc + "output_conv2.0.weight"], sd[sc + "output_conv2.0.bias"], padding=1)
        o = F.relu(o)
        o = F.conv2d(o, sd[sc + "output_conv2.2.weight"], sd[sc + "output_conv2.2.bias"])
        o = F.relu(o)
    return o.to(ori)


def head_single_image(sd, taps, ph: int, pw: int, T: int, motion_features, pred_depth_idx=None,
                      skip_tmp_block: bool = False):
    """`DPTHeadTemporal.foward_single_image`, dpt_temporal.py:132-260.  taps: the new frame's 4 encoder taps [1, N, D];
    motion_features: stored (layer_1..4) of the T-1 context frames, NCHW.  Only the frames `pred_depth_idx` + the new
    one go through refinenet2/1 and the output convs (the reference's chunks of `max_head_pred_length` frames are a
    memory measure: every op of that part is per frame).  Returns (depth [n,1,H,W], layer_1..4 of the new frame)."""
    h, sc = "head.", "head.scratch."
    l1, l2, l3, l4 = reassemble(sd, taps, ph, pw)
    o1, o2, o3, o4 = motion_features
    if pred_depth_idx is not None:
        l1p, l2p = torch.cat([o1[pred_depth_idx], l1], 0), torch.cat([o2[pred_depth_idx], l2], 0)
    else:
        l1p, l2p = l1, l2
    l3c, l4c = torch.cat([o3, l3], 0), torch.cat([o4, l4], 0)
    l1r = F.conv2d(l1p, sd[sc + "layer1_rn.weight"], None, padding=1)
    l2r = F.conv2d(l2p, sd[sc + "layer2_rn.weight"], None, padding=1)
    l4r = F.conv2d(temporal_module(sd, h + "motion_modules.1.", l4c, T), sd[sc + "layer4_rn.weight"], None, padding=1)
    l3r = F.conv2d(temporal_module(sd, h + "motion_modules.0.", l3c, T), sd[sc + "layer3_rn.weight"], None, padding=1)
    p4 = _fusion(sd, sc + "refinenet4.", l4r, None, l3r.shape[2:])
    if not skip_tmp_block:
        p4 = temporal_module(sd, h + "motion_modules.2.", p4, T)
    p3 = _fusion(sd, sc + "refinenet3.", p4, l3r, l2r.shape[2:])
    p3 = temporal_module(sd, h + "motion_modules.3.", p3, T)
    p3 = p3[list(pred_depth_idx) + [p3.shape[0] - 1]] if pred_depth_idx is not None else p3[-1:]
    p2 = _fusion(sd, sc + "refinenet2.", p3, l2r, l1r.shape[2:])
    p1 = _fusion(sd, sc + "refinenet1.", p2, l1r, None)
    return _output_convs(sd, p1, ph, pw), l1, l2, l3, l4


def forward_single_image(sd, x: torch.Tensor, encoder: str, motion_features, pred_depth_idx=None, inference_length=32,
                         skip_tmp_block: bool = False):
    """`VideoDepthAnything.forward_single_image`, video_depth.py:67-91.  x [1,1,3,H,W]."""
    B, T, _, H, W = x.shape
    ph, pw = H // PATCH, W // PATCH
    taps = encoder_taps(sd, x.flatten(0, 1), encoder)
    d, l1, l2, l3, l4 = head_single_image(sd, taps, ph, pw, inference_length, motion_features, pred_depth_idx, skip_tmp_block)
    d = F.relu(F.interpolate(d, size=(H, W), mode="bilinear", align_corners=True))
    n = T if pred_depth_idx is None else len(pred_depth_idx) + 1
    return d.squeeze(1).unflatten(0, (B, n)), (l1, l2, l3, l4)


def head_forward(sd, taps, ph: int, pw: int, T: int, skip_tmp_block: bool = False, collect=None) -> torch.Tensor:
    """`DPTHeadTemporal.forward`, dpt_temporal.py:53-99 (layers built in dpt.py:47-124)."""
    h = "head."
    layers = []
    for i, x in enumerate(taps):
        BT, _, D = x.shape
        x = x.permute(0, 2, 1).reshape(BT, D, ph, pw)
        x = F.conv2d(x, sd[f"{h}projects.{i}.weight"], sd[f"{h}projects.{i}.bias"])
        if i == 0:
            x = F.conv_transpose2d(x, sd[h + "resize_layers.0.weight"], sd[h + "resize_layers.0.bias"], stride=4)
        elif i == 1:
            x = F.conv_transpose2d(x, sd[h + "resize_layers.1.weight"], sd[h + "resize_layers.1.bias"], stride=2)
        elif i == 3:
            x = F.conv2d(x, sd[h + "resize_layers.3.weight"], sd[h + "resize_layers.3.bias"], stride=2, padding=1)
        layers.append(x)
    l1, l2, l3, l4 = layers
    if collect is not None:
        for i, x in enumerate(layers):
            collect[f"layer_{i + 1}"] = x
    l3 = temporal_module(sd, h + "motion_modules.0.", l3, T)
    l4 = temporal_module(sd, h + "motion_modules.1.", l4, T)
    if collect is not None:
        collect["mm0"], collect["mm1"] = l3, l4
    sc = h + "scratch."
    l1r = F.conv2d(l1, sd[sc + "layer1_rn.weight"], None, padding=1)
    l2r = F.conv2d(l2, sd[sc + "layer2_rn.weight"], None, padding=1)
    l3r = F.conv2d(l3, sd[sc + "layer3_rn.weight"], None, padding=1)
    l4r = F.conv2d(l4, sd[sc + "layer4_rn.weight"], None, padding=1)
    p4 = _fusion(sd, sc + "refinenet4.", l4r, None, l3r.shape[2:])
    if collect is not None:
        collect["l1r"], collect["l2r"], collect["l3r"], collect["l4r"] = l1r, l2r, l3r, l4r
        collect["path_4_pre"] = p4   # refinenet4's output, before motion module 2
    if not skip_tmp_block:
        p4 = temporal_module(sd, h + "motion_modules.2.", p4, T)
    p3 = _fusion(sd, sc + "refinenet3.", p4, l3r, l2r.shape[2:])
    if collect is not None:
        collect["path_3_pre"] = p3   # refinenet3's output, before motion module 3
    p3 = temporal_module(sd, h + "motion_modules.3.", p3, T)
    p2 = _fusion(sd, sc + "refinenet2.", p3, l2r, l1r.shape[2:])
    p1 = _fusion(sd, sc + "refinenet1.", p2, l1r, None)
    if collect is not None:
        collect["path_4"], collect["path_3"], collect["path_2"], collect["path_1"] = p4, p3, p2, p1
    out = F.conv2d(p1, sd[sc + "output_conv1.weight"], sd[sc + "output_conv1.bias"], padding=1)
    out = F.interpolate(out, (ph * PATCH, pw * PATCH), mode="bilinear", align_corners=True)
    ori = out.dtype
    with torch.autocast(device_type="cuda", enabled=False):  # the fp32 island, dpt_temporal.py:95-97
        o = F.conv2d(out.float(), sd[sc + "output_conv2.0.weight"], sd[sc + "output_conv2.0.bias"], padding=1)
        o = F.relu(o)
        o = F.conv2d(o, sd[sc + "output_conv2.2.weight"], sd[sc + "output_conv2.2.bias"])
        o = F.relu(o)
    return o.to(ori)


def forward(sd, x: torch.Tensor, encoder: str, skip_tmp_block: bool = False, collect=None) -> torch.Tensor:
    """`VideoDepthAnything.forward`, video_depth.py:58-65.  x [B,T,3,H,W] -> [B,T,H,W]."""
    assert encoder in MODEL_CONFIGS
    B, T, _, H, W = x.shape
    ph, pw = H // PATCH, W // PATCH
    taps = encoder_taps(sd, x.flatten(0, 1), encoder, collect)
    depth = head_forward(sd, taps, ph, pw, T, skip_tmp_block, collect)
    depth = F.interpolate(depth, size=(H, W), mode="bilinear", align_corners=True)
    depth = F.relu(depth)
    return depth.squeeze(1).unflatten(0, (B, T))

"""Parameter containers that reproduce the reference's module tree -- and therefore its state-dict key set,
shapes and default initialisation -- without any of its compute.  None of these modules' `forward` is ever
called: compute goes through `engine.ForwardPlan` into libvda_sm100a.so.

Key layout mirrored from the reference (file:line):
  pretrained.*            video_depth_anything/dinov2.py:98-178 (cls_token, pos_embed, mask_token, patch_embed, blocks, norm)
  pretrained.blocks.N.*   dinov2_layers/block.py:37-81, attention.py:29-47, mlp.py:17-33, layer_scale.py:15-25
  head.projects/resize_layers/scratch.*   dpt.py:47-124, util/blocks.py:4-35,38-66,94-133
  head.motion_modules.N.* dpt_temporal.py:35-51, motion_module/motion_module.py:74-107,136-171,189-246, attention.py:31-91,296-384
"""
from __future__ import annotations

import math
import weakref

import torch
import torch.nn as nn

ENCODERS = {
    # dinov2.py:339-378 ; taps video_depth.py:48-51
    "vits": dict(dim=384, depth=12, heads=6, taps=[2, 5, 8, 11]),
    "vitl": dict(dim=1024, depth=24, heads=16, taps=[4, 11, 17, 23]),
}


class _Attn(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.qkv = nn.Linear(dim, dim * 3, bias=True)
        self.proj = nn.Linear(dim, dim, bias=True)


class _Mlp(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.fc1 = nn.Linear(dim, 4 * dim)
        self.fc2 = nn.Linear(4 * dim, dim)


class _LayerScale(nn.Module):
    def __init__(self, dim, init_values=1.0):
        super().__init__()
        self.gamma = nn.Parameter(init_values * torch.ones(dim))


class _Block(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=1e-6)
        self.attn = _Attn(dim)
        self.ls1 = _LayerScale(dim)
        self.norm2 = nn.LayerNorm(dim, eps=1e-6)
        self.mlp = _Mlp(dim)
        self.ls2 = _LayerScale(dim)


class _PatchEmbed(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.proj = nn.Conv2d(3, dim, kernel_size=14, stride=14)


class DinoV2Params(nn.Module):
    """`DinoVisionTransformer` parameter tree (img 518, patch 14, no register tokens; dinov2.py:395-415)."""

    def __init__(self, encoder: str):
        super().__init__()
        cfg = ENCODERS[encoder]
        D = cfg["dim"]
        self.embed_dim = D
        self.num_heads = cfg["heads"]
        self.n_blocks = cfg["depth"]
        self.patch_size = 14
        self.cls_token = nn.Parameter(torch.zeros(1, 1, D))
        self.pos_embed = nn.Parameter(torch.zeros(1, 37 * 37 + 1, D))
        self.mask_token = nn.Parameter(torch.zeros(1, D))
        self.patch_embed = _PatchEmbed(D)
        self.blocks = nn.ModuleList([_Block(D) for _ in range(cfg["depth"])])
        self.norm = nn.LayerNorm(D, eps=1e-6)
        nn.init.trunc_normal_(self.pos_embed, std=0.02)  # dinov2.py:172-177
        nn.init.normal_(self.cls_token, std=1e-6)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                nn.init.zeros_(m.bias)


class _RCU(nn.Module):
    def __init__(self, f):
        super().__init__()
        self.conv1 = nn.Conv2d(f, f, 3, 1, 1, bias=True)
        self.conv2 = nn.Conv2d(f, f, 3, 1, 1, bias=True)


class _Fusion(nn.Module):
    def __init__(self, f):
        super().__init__()
        self.out_conv = nn.Conv2d(f, f, 1, 1, 0, bias=True)
        self.resConfUnit1 = _RCU(f)
        self.resConfUnit2 = _RCU(f)


class _Scratch(nn.Module):
    def __init__(self, oc, f):
        super().__init__()
        self.layer1_rn = nn.Conv2d(oc[0], f, 3, 1, 1, bias=False)
        self.layer2_rn = nn.Conv2d(oc[1], f, 3, 1, 1, bias=False)
        self.layer3_rn = nn.Conv2d(oc[2], f, 3, 1, 1, bias=False)
        self.layer4_rn = nn.Conv2d(oc[3], f, 3, 1, 1, bias=False)
        self.refinenet1 = _Fusion(f)
        self.refinenet2 = _Fusion(f)
        self.refinenet3 = _Fusion(f)
        self.refinenet4 = _Fusion(f)
        self.output_conv1 = nn.Conv2d(f, f // 2, 3, 1, 1)
        self.output_conv2 = nn.Sequential(
            nn.Conv2d(f // 2, 32, 3, 1, 1), nn.ReLU(True), nn.Conv2d(32, 1, 1, 1, 0), nn.ReLU(True), nn.Identity()
        )


class _PosEnc(nn.Module):
    def __init__(self, d_model, max_len):
        super().__init__()
        position = torch.arange(max_len).unsqueeze(1)
        div_term = torch.exp(torch.arange(0, d_model, 2) * (-math.log(10000.0) / d_model))
        pe = torch.zeros(1, max_len, d_model)
        pe[0, :, 0::2] = torch.sin(position * div_term)
        pe[0, :, 1::2] = torch.cos(position * div_term)
        self.register_buffer("pe", pe)


class _TemporalAttention(nn.Module):
    def __init__(self, dim, max_len, pe="ape"):
        super().__init__()
        self.to_q = nn.Linear(dim, dim, bias=False)
        self.to_k = nn.Linear(dim, dim, bias=False)
        self.to_v = nn.Linear(dim, dim, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(dim, dim), nn.Dropout(0.0)])
        # 'ape': a `pe` buffer in the state dict; 'rope': freqs_cis is a plain attribute, no state-dict entry
        # (motion_module.py:226-240)
        self.pos_embedding_type = pe
        self.temporal_max_len = max_len
        self.pos_encoder = _PosEnc(dim, max_len) if pe == "ape" else None


class _GEGLU(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.proj = nn.Linear(dim, dim * 8)


class _FeedForward(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.net = nn.ModuleList([_GEGLU(dim), nn.Dropout(0.0), nn.Linear(dim * 4, dim)])


class _TemporalBlock(nn.Module):
    def __init__(self, dim, max_len, pe="ape"):
        super().__init__()
        self.attention_blocks = nn.ModuleList([_TemporalAttention(dim, max_len, pe) for _ in range(2)])
        self.norms = nn.ModuleList([nn.LayerNorm(dim) for _ in range(2)])
        self.ff = _FeedForward(dim)
        self.ff_norm = nn.LayerNorm(dim)


class _TemporalTransformer(nn.Module):
    def __init__(self, C, max_len, pe="ape"):
        super().__init__()
        self.norm = nn.GroupNorm(32, C, eps=1e-6, affine=True)
        self.proj_in = nn.Linear(C, C)
        self.transformer_blocks = nn.ModuleList([_TemporalBlock(C, max_len, pe)])
        self.proj_out = nn.Linear(C, C)
        for p in self.proj_out.parameters():  # zero_module, motion_module.py:61-62
            p.detach().zero_()


class TemporalModuleParams(nn.Module):
    def __init__(self, C, max_len, pe="ape"):
        super().__init__()
        self.temporal_transformer = _TemporalTransformer(C, max_len, pe)


class DPTHeadTemporalParams(nn.Module):
    """`DPTHeadTemporal` parameter tree (use_bn=False, use_clstoken=False: the only configuration any caller uses)."""

    def __init__(self, in_channels, features, out_channels, num_frames, pe="ape"):
        super().__init__()
        oc = out_channels
        self.projects = nn.ModuleList([nn.Conv2d(in_channels, c, 1, 1, 0) for c in oc])
        self.resize_layers = nn.ModuleList([
            nn.ConvTranspose2d(oc[0], oc[0], 4, 4, 0),
            nn.ConvTranspose2d(oc[1], oc[1], 2, 2, 0),
            nn.Identity(),
            nn.Conv2d(oc[3], oc[3], 3, 2, 1),
        ])
        self.scratch = _Scratch(oc, features)
        self.motion_modules = nn.ModuleList(
            [TemporalModuleParams(c, num_frames, pe) for c in (oc[2], oc[3], features, features)]
        )

    def _bind(self, owner):
        """The compute lives in the owning model's launch plans; a weak reference keeps the module tree acyclic."""
        object.__setattr__(self, "_owner", weakref.ref(owner))

    def __getstate__(self):   # copy / pickle: the owner re-binds its head (VideoDepthAnything.__setstate__)
        state = self.__dict__.copy()
        state.pop("_owner", None)
        return state

    def _model(self):
        owner = getattr(self, "_owner", None)
        owner = owner() if owner is not None else None
        if owner is None:
            raise RuntimeError("this head is not attached to a VideoDepthAnything model")
        return owner

    def get_motion_features(self, out_features, patch_h, patch_w):
        """projects + resize layers of n frames' encoder features (dpt_temporal.py:101-130): out_features = four
        (patch tokens [n, P, D], class token) pairs as `get_intermediate_layers(..., return_class_token=True)` returns
        them.  Returns layer_1..4, NCHW."""
        model = self._model()
        tok0 = out_features[0][0] if isinstance(out_features[0], (tuple, list)) else out_features[0]
        n, dev = tok0.shape[0], tok0.device
        bf16 = torch.is_autocast_enabled("cuda")
        outs = [[] for _ in range(4)]
        with torch.cuda.device(dev), torch.no_grad():
            plan = model.stream_plan(1, 14 * patch_h, 14 * patch_w, dev, bf16, 1, (1,), False, enc_batch=1)
            for i in range(n):
                for tap, f in zip(plan.enc_taps, out_features):
                    tok = f[0] if isinstance(f, (tuple, list)) else f
                    tap.copy_(tok[i].reshape(patch_h * patch_w, -1))
                plan.run_stage("reassemble")
                for k in range(4):
                    outs[k].append(plan.load_features(k, 1))
        return tuple(torch.cat(o, 0) for o in outs)

    def foward_single_image(self, out_features, patch_h, patch_w, frame_length, motion_features, pred_depth_idx=None,
                            skip_tmp_block=False, max_head_pred_length=6):
        """The head's streaming step (dpt_temporal.py:133-260; the reference spells it `foward`): the new frame's encoder
        features + the stored layer_1..4 of the frame_length - 1 context frames -> (depth [n, 1, H, W], layer_1..4 of the
        new frame).  `max_head_pred_length` only chunks the reference's per-frame tail to save memory; ignored."""
        depth, feats = self._model()._single_image(None, out_features, patch_h, patch_w, frame_length, motion_features,
                                                   pred_depth_idx, skip_tmp_block)
        return (depth,) + tuple(feats)

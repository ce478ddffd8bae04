"""Drop-in for `video_depth_anything.video_depth.VideoDepthAnything` (reference video_depth.py:36-65,329-417).

Same constructor, same state-dict key set (so `load_state_dict(torch.load(ckpt), strict=True)` works unchanged),
same `forward(x, skip_tmp_block=False)` and `infer_video_depth(frames, target_fps, input_size=518, device='cuda',
fp32=False, skip_tmp_block=False)` signatures and return types.  All compute runs in libvda_sm100a.so (hand-written
sm_100a kernels); there is no eager / CPU fallback: a missing library or a non-sm_100 device raises.

Precision follows the reference's contract: under CUDA autocast (what `infer_video_depth(fp32=False)` enters) the
bf16 tensor-core build runs and reproduces autocast's rounding points; without autocast the fp32 build runs.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from . import lib as L
from .driver import WindowedDepthDriver
from .engine import ForwardPlan, HeadPlan, PackedWeights, StreamPlan
from .modules import DinoV2Params, DPTHeadTemporalParams

# infer settings, do not change (video_depth.py:29-33); imported by the reference's other files
INFER_LEN = 32
OVERLAP = 10
KEYFRAMES = [0, 12, 24, 25, 26, 27, 28, 29, 30, 31]
INTERP_LEN = 8


class VideoDepthAnything(nn.Module):
    def __init__(self, encoder="vitl", features=256, out_channels=[256, 512, 1024, 1024], use_bn=False,
                 use_clstoken=False, num_frames=32, pe="ape"):
        super().__init__()
        if use_bn or use_clstoken:
            raise NotImplementedError("use_bn / use_clstoken are never enabled by the reference's callers (run.py:74-77)")
        if pe not in ("ape", "rope"):
            raise NotImplementedError(pe)  # motion_module.py:242-243
        self.intermediate_layer_idx = {"vits": [2, 5, 8, 11], "vitl": [4, 11, 17, 23]}
        if encoder not in self.intermediate_layer_idx:
            raise KeyError(encoder)
        self.encoder = encoder
        self.num_frames = num_frames
        self.pe = pe
        self.pretrained = DinoV2Params(encoder)
        self.head = DPTHeadTemporalParams(self.pretrained.embed_dim, features, list(out_channels), num_frames, pe)
        self._packed = {}   # (device, bf16) -> (version key, PackedWeights)
        self._plans = {}    # (device, bf16, B, T, H, W, skip) -> ForwardPlan
        self.force_simt = False  # tests: run the bf16 build on the SIMT kernels instead of tcgen05
        self.attn_simt = False   # tests: tcgen05 GEMMs but the exact-rounding SIMT attention

    # ------------------------------------------------------------------ weights / plans
    def _weights_version(self):
        return tuple((p.data_ptr(), p._version) for p in self.parameters())

    def _packed_weights(self, device, bf16):
        key = (str(device), bf16)
        ver = self._weights_version()
        hit = self._packed.get(key)
        if hit is None or hit[0] != ver:
            self._plans = {k: v for k, v in self._plans.items() if (k[0], k[1]) != key}
            self._packed[key] = (ver, PackedWeights(self, device, bf16))
        return self._packed[key][1]

    def plan(self, B, T, H, W, device, bf16, skip_tmp_block=False, keep=()):
        pw = self._packed_weights(device, bf16)
        use_tc = bf16 and not self.force_simt
        key = (str(device), bf16, B, T, H, W, bool(skip_tmp_block), use_tc, self.attn_simt, tuple(sorted(keep)))
        p = self._plans.get(key)
        if p is None:
            p = ForwardPlan(pw, B, T, H, W, device, skip_tmp_block, use_tc=use_tc, keep=keep, attn_tc=use_tc and not self.attn_simt)
            self._plans[key] = p
        return p

    def head_plan(self, B, T, H, W, device, bf16, skip_tmp_block=False, keep=()):
        """Plan of the head alone (taps in, depth out) for x of [B, T, 3, H, W]."""
        pw = self._packed_weights(device, bf16)
        use_tc = bf16 and not self.force_simt
        key = (str(device), bf16, "head", B, T, H, W, bool(skip_tmp_block), use_tc, tuple(sorted(keep)))
        p = self._plans.get(key)
        if p is None:
            p = HeadPlan(pw, B, T, H, W, device, skip_tmp_block, use_tc=use_tc, keep=keep, attn_tc=use_tc)
            self._plans[key] = p
        return p

    def stream_plan(self, T, H, W, device, bf16, rows, npreds, skip_tmp_block=False, enc_batch=1):
        """Plan of the streaming mode for one (context length, network size, store size, prediction counts)."""
        pw = self._packed_weights(device, bf16)
        key = (str(device), bf16, "stream", T, H, W, int(rows), tuple(npreds), bool(skip_tmp_block), self.force_simt, self.attn_simt,
               int(enc_batch))
        p = self._plans.get(key)
        if p is None:
            p = StreamPlan(pw, T, H, W, device, rows, npreds, skip_tmp_block, use_tc=bf16 and not self.force_simt,
                           attn_tc=bf16 and not self.force_simt and not self.attn_simt, enc_batch=enc_batch)
            self._plans[key] = p
        return p

    # ------------------------------------------------------------------ forward (video_depth.py:58-65)
    def forward(self, x, skip_tmp_block=False):
        B, T, Cc, H, W = x.shape
        L.require_device(x.device)
        bf16 = torch.is_autocast_enabled("cuda") if hasattr(torch, "is_autocast_enabled") else False
        plan = self.plan(B, T, H, W, x.device, bf16, skip_tmp_block)
        depth = plan.run(x.to(torch.float32).flatten(0, 1))
        # fp32 in both builds: under CUDA autocast the head's bilinear upsample is an fp32 op, so the output convs and
        # the result stay fp32; the identity resize + ReLU of video_depth.py:63-64 are numerical no-ops.
        return depth.clone().unflatten(0, (B, T))

    # ------------------------------------------------------------------ streaming mode (video_depth.py:93-327)
    def infere_single_image(self, frames, target_fps, input_size=518, device="cuda", fp32=False, warmup=True,
                            inference_length=32, keyframe_list=[0, 12], align_each_new_frame=True, skip_tmp_block=False):
        """One new frame per step against a stored temporal context (the fork's streaming mode).  Same arguments, return
        value and failure modes as the reference: `warmup=False` raises NotImplementedError (video_depth.py:321-322); with
        `align_each_new_frame` the first frame only serves the alignment and is dropped from the result (:324-325), and
        keyframe lists whose rows fall outside the first context raise IndexError as the reference does (its default
        `[0, 12]` among them -- use keyframes >= 2, or align_each_new_frame=False)."""
        if not warmup:
            raise NotImplementedError
        from .stream import StreamingDepthDriver
        drv = StreamingDepthDriver(self, device=device, fp32=fp32, skip_tmp_block=skip_tmp_block)
        depth = drv.run(frames, input_size, inference_length, keyframe_list, align_each_new_frame)
        return depth, target_fps

    # ------------------------------------------------------------------ driver (video_depth.py:329-417)
    def infer_video_depth(self, frames, target_fps, input_size=518, device="cuda", fp32=False, skip_tmp_block=False):
        drv = WindowedDepthDriver(self, device=device, fp32=fp32, skip_tmp_block=skip_tmp_block)
        depth = drv.run(frames, input_size)
        return depth, target_fps

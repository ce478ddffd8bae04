"""Drop-in for `video_depth_anything.video_depth.VideoDepthAnything` (reference video_depth.py:36-65,329-417).

Same constructor, same state-dict key set (so `load_state_dict(torch.load(ckpt), strict=True)` works unchanged),
same `forward(x, skip_tmp_block=False)` and `infer_video_depth(frames, target_fps, input_size=518, device='cuda',
fp32=False, skip_tmp_block=False)` signatures and return types.  All compute runs in libvda_sm100a.so (hand-written
sm_100a kernels); there is no eager / CPU fallback: a missing library or a non-sm_100 device raises.

Precision follows the reference's contract: under CUDA autocast (what `infer_video_depth(fp32=False)` enters) the
bf16 tensor-core build runs and reproduces autocast's rounding points; without autocast the fp32 build runs.
"""
from __future__ import annotations

import collections
import warnings

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
        self.head._bind(self)
        self._packed = {}   # (device, bf16) -> (version key, PackedWeights)
        # (device, bf16, B, T, H, W, skip, ...) -> plan, least recently used first.  A plan owns every activation buffer of
        # its shape (vitl 518x518: ~20 GB), so the cache is capped: callers that walk over many resolutions (app.py,
        # eval loops) evict the oldest plan instead of exhausting the GPU.
        self._plans = collections.OrderedDict()
        self.max_plans = 4
        self._fp16_autocast_noted = False
        self.force_simt = False  # tests: run the bf16 build on the SIMT kernels instead of tcgen05
        self.attn_simt = False   # tests: tcgen05 GEMMs but the exact-rounding SIMT attention

    def __getstate__(self):   # copy.deepcopy / torch.save(model): launch plans and packed weights are rebuilt on demand
        state = self.__dict__.copy()
        state["_packed"], state["_plans"] = {}, collections.OrderedDict()
        state.pop("_last_driver", None)
        return state

    def __setstate__(self, state):
        super().__setstate__(state)
        self.head._bind(self)

    # ------------------------------------------------------------------ weights / plans
    def _weights_version(self):
        return tuple((p.data_ptr(), p._version) for p in self.parameters())

    def _packed_weights(self, device, bf16):
        key = (str(device), bf16)
        ver = self._weights_version()
        hit = self._packed.get(key)
        if hit is None or hit[0] != ver:
            self._plans = collections.OrderedDict((k, v) for k, v in self._plans.items() if (k[0], k[1]) != key)
            self._packed[key] = (ver, PackedWeights(self, device, bf16))
        return self._packed[key][1]

    def _cached_plan(self, key, build):
        p = self._plans.get(key)
        if p is None:
            while len(self._plans) >= max(1, int(self.max_plans)):
                self._plans.popitem(last=False)   # drop the least recently used plan; its buffers go back to the allocator
            p = build()
            self._plans[key] = p
        else:
            self._plans.move_to_end(key)
        return p

    def plan(self, B, T, H, W, device, bf16, skip_tmp_block=False, keep=()):
        pw = self._packed_weights(device, bf16)
        use_tc = bf16 and not self.force_simt
        key = (str(device), bf16, B, T, H, W, bool(skip_tmp_block), use_tc, self.attn_simt, tuple(sorted(keep)))
        return self._cached_plan(key, lambda: ForwardPlan(pw, B, T, H, W, device, skip_tmp_block, use_tc=use_tc, keep=keep,
                                                          attn_tc=use_tc and not self.attn_simt))

    def head_plan(self, B, T, H, W, device, bf16, skip_tmp_block=False, keep=()):
        """Plan of the head alone (taps in, depth out) for x of [B, T, 3, H, W]."""
        pw = self._packed_weights(device, bf16)
        use_tc = bf16 and not self.force_simt
        key = (str(device), bf16, "head", B, T, H, W, bool(skip_tmp_block), use_tc, tuple(sorted(keep)))
        return self._cached_plan(key, lambda: HeadPlan(pw, B, T, H, W, device, skip_tmp_block, use_tc=use_tc, keep=keep, attn_tc=use_tc))

    def stream_plan(self, T, H, W, device, bf16, rows, npreds, skip_tmp_block=False, enc_batch=1):
        """Plan of the streaming mode for one (context length, network size, store size, prediction counts)."""
        pw = self._packed_weights(device, bf16)
        key = (str(device), bf16, "stream", T, H, W, int(rows), tuple(npreds), bool(skip_tmp_block), self.force_simt, self.attn_simt,
               int(enc_batch))
        return self._cached_plan(key, lambda: StreamPlan(pw, T, H, W, device, rows, npreds, skip_tmp_block,
                                                         use_tc=bf16 and not self.force_simt,
                                                         attn_tc=bf16 and not self.force_simt and not self.attn_simt,
                                                         enc_batch=enc_batch))

    # ------------------------------------------------------------------ forward (video_depth.py:58-65)
    def forward(self, x, skip_tmp_block=False):
        B, T, Cc, H, W = x.shape
        L.require_device(x.device)
        bf16 = torch.is_autocast_enabled("cuda")
        if bf16 and torch.get_autocast_dtype("cuda") == torch.float16 and not self._fp16_autocast_noted:
            # the reference's `infer_video_depth(fp32=False)` enters a dtype-less autocast, i.e. fp16 on CUDA
            # (video_depth.py:367); this library has one reduced-precision build, bf16 (BASELINE.json north_star)
            self._fp16_autocast_noted = True
            warnings.warn("video_depth_anything_b200: CUDA autocast dtype is float16; the bf16 build runs instead "
                          "(wider exponent, 3 fewer mantissa bits) -- see INTEGRATION.md", RuntimeWarning, stacklevel=2)
        plan = self.plan(B, T, H, W, x.device, bf16, skip_tmp_block)
        depth = plan.run(x.to(torch.float32).flatten(0, 1))
        self._fire_motion_module_hooks(plan, B, T, skip_tmp_block)
        # fp32 in both builds: under CUDA autocast the head's bilinear upsample is an fp32 op, so the output convs and
        # the result stay fp32; the identity resize + ReLU of video_depth.py:63-64 are numerical no-ops.
        return depth.clone().unflatten(0, (B, T))

    def _fire_motion_module_hooks(self, plan, B, T, skip_tmp_block):
        """Forward hooks registered on `model.head.motion_modules[i]` -- the reference's diagnostic harness grabs the temporal
        modules' outputs that way (Test_singleImage_processing.py:132-142).  The modules here are parameter containers whose
        `forward` never runs (the compute is a launch list), so the hooks are called by hand after the plan has run, with the
        tensors the reference would pass: hook(module, (input [B,C,T,h,w], None, None), output [B,C,T,h,w]) as at
        dpt_temporal.py:75-76,84-87 -- copies of the plan's buffers, in the build's activation type."""
        mods = self.head.motion_modules
        if not any(m._forward_hooks for m in mods):
            return
        ph, pw = plan.ph, plan.pw_
        geo = [(ph, pw), ((ph + 1) // 2, (pw + 1) // 2), (ph, pw), (2 * ph, 2 * pw)]
        names = [("layer_3", "mm0"), ("layer_4", "mm1"), ("path_4d", "path_4"), ("path_3d", "path_3")]
        for i, m in enumerate(mods):
            if not m._forward_hooks or (i == 2 and skip_tmp_block):
                continue
            h, w = geo[i]
            xin = plan.bufs[names[i][0]].view(B, T, h, w, -1).permute(0, 4, 1, 2, 3).contiguous()
            out = plan.bufs[names[i][1]].view(B, T, h + 2, w + 2, -1)[:, :, 1:-1, 1:-1].permute(0, 4, 1, 2, 3).contiguous()
            for hook in list(m._forward_hooks.values()):
                if hook(m, (xin, None, None), out) is not None:
                    raise NotImplementedError("forward hooks on motion modules can observe the output but not replace it")

    # ------------------------------------------------------------------ tensor-level streaming step (video_depth.py:67-91)
    def forward_single_image(self, x, motion_features, pred_depth_idx=None, inference_length=32, skip_tmp_block=False):
        """One new frame against caller-held context features: same arguments and return value as the reference.
        x [1, 1, 3, H, W]; motion_features = (layer_1, layer_2, layer_3, layer_4) of the `inference_length - 1` context
        frames, NCHW as `head.get_motion_features` / an earlier call returned them; pred_depth_idx: context rows that get
        a depth map besides the new frame.  Returns (depth [1, n, H, W] fp32, (layer_1..4 of the new frame, [1, C, h, w])).

        This is the reference's calling convention kept for callers that manage the store themselves: the four context
        tensors are converted into the plan's channels-last store on every call.  `infere_single_image` keeps the store
        on the device in that layout and is the fast path."""
        B, T, _, H, W = x.shape
        if B * T != 1:
            raise ValueError(f"forward_single_image takes one frame [1, 1, 3, H, W], got {list(x.shape)}")
        depth, feats = self._single_image(x, None, H // 14, W // 14, inference_length, motion_features, pred_depth_idx,
                                          skip_tmp_block)
        return depth.squeeze(1).unflatten(0, (B, depth.shape[0])), feats

    def _single_image(self, x, out_features, patch_h, patch_w, frame_length, motion_features, pred_depth_idx, skip_tmp_block):
        """Shared by `forward_single_image` (x given: the encoder runs) and `head.foward_single_image` (encoder features
        given).  Returns (depth [n, 1, H, W] fp32, (layer_1..4 of the new frame))."""
        ref = x if x is not None else out_features[0][0]
        L.require_device(ref.device)
        dev = ref.device
        bf16 = torch.is_autocast_enabled("cuda")
        T = int(frame_length)
        H, W = 14 * patch_h, 14 * patch_w
        if len(motion_features) != 4:
            raise ValueError("motion_features must be (layer_1, layer_2, layer_3, layer_4)")
        for k, t in enumerate(motion_features):
            if t.shape[0] != T - 1:
                # the reference fails in unflatten(0, (B, T)) when context + new frame is not exactly frame_length frames
                raise RuntimeError(f"motion_features[{k}] holds {t.shape[0]} frames, frame_length - 1 = {T - 1} expected")
        pred = [] if pred_depth_idx is None else [int(i) % (T - 1) if i < 0 else int(i) for i in pred_depth_idx]
        if any(i >= T - 1 for i in pred):
            raise IndexError(f"pred_depth_idx {pred_depth_idx} out of range for {T - 1} context frames")
        n = len(pred) + 1
        R = T - 1   # store rows 0..T-2 hold the context, slot R receives the new frame
        with torch.cuda.device(dev), torch.no_grad():
            plan = self.stream_plan(T, H, W, dev, bf16, R, (n,), skip_tmp_block, enc_batch=1)
            for k, t in enumerate(motion_features):
                plan.store_features(k, 0, t)
            plan.load_indices(plan.index_row(ctx=list(range(T - 1)) + [R], pred=pred + [R], p3=pred + [T - 1]).to(dev))
            if x is not None:
                plan.x_in.copy_(x.reshape(1, 3, H, W).to(torch.float32))
                plan.run_stage("encode")
            else:
                if len(out_features) != 4:
                    raise ValueError("out_features must hold the four encoder taps")
                P = patch_h * patch_w
                for tap, f in zip(plan.enc_taps, out_features):
                    tok = f[0] if isinstance(f, (tuple, list)) else f   # (patch tokens, class token), dinov2.py:309-321
                    if tuple(tok.shape) != (1, P, tap.shape[1]):
                        raise ValueError(f"encoder feature {list(tok.shape)}: expected [1, {P}, {tap.shape[1]}]")
                    tap.copy_(tok.reshape(P, -1))
                plan.run_stage("reassemble")
            plan.run_stage("middle")
            plan.run_stage(f"tail{n}")
            depth = plan.depths[n].reshape(n, 1, H, W).clone()
            feats = tuple(plan.load_features(k, R) for k in range(4))
        return depth, feats

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
        """Same arguments and return value as the reference.  Runs in this process only, whatever torch.distributed state
        exists (one video per rank under torchrun behaves like the reference); `infer_video_depth_sharded` splits ONE video
        over the ranks of a process group."""
        drv = WindowedDepthDriver(self, device=device, fp32=fp32, skip_tmp_block=skip_tmp_block)
        depth = drv.run(frames, input_size)
        return depth, target_fps

    def infer_video_depth_sharded(self, frames, target_fps, input_size=518, device="cuda", fp32=False, skip_tmp_block=False,
                                  group=None, return_device=False):
        """One video over the GPUs of a node: every rank of `group` (default: the world group; one
        process per GPU, same node) calls this with the same `frames`, of which it reads only the frames its own windows
        touch.  Window k runs on rank k % world; the ranks exchange 11 frames per window over NCCL, each stitches its own
        windows (bit-identical to a single-rank run) and copies them to a host buffer shared by the ranks.

        Returns (depth, target_fps): on group rank 0 `depth` is the [N, H, W] fp32 result as a numpy view of that shared
        page-locked buffer -- valid until the next sharded call on this model, copy it to keep it; on the other ranks it
        is None.  With return_device nothing is copied to the host: every rank gets its own final frames as a device
        tensor (`WindowedDepthDriver.last_frame_ranges` names them).  The ranks must agree on the frame count and size
        (checked)."""
        dist = torch.distributed
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("infer_video_depth_sharded needs an initialised torch.distributed process group")
        group = dist.group.WORLD if group is None else group
        drv = WindowedDepthDriver(self, device=device, fp32=fp32, skip_tmp_block=skip_tmp_block, group=group)
        # the exchange sizes follow from the video geometry: refuse to start (instead of hanging in a collective) when the
        # ranks were handed different videos
        mine = [int(frames.shape[0])] + [int(v) for v in frames[0].shape[:2]] + [int(input_size)]
        seen = [None] * dist.get_world_size(group)
        dist.all_gather_object(seen, mine, group=group)
        if any(v != mine for v in seen):
            raise ValueError(f"infer_video_depth_sharded: ranks disagree on (frames, H, W, input_size): {seen}")
        self._last_driver = drv
        depth = drv.run(frames, input_size, return_device=return_device)
        return depth, target_fps

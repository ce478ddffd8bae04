"""Host side of the forward pass: weight packing and the per-shape launch plan.

`PackedWeights` turns the reference-layout state dict into the K-major matrices the kernels consume (once per
device / precision).  `ForwardPlan` owns every activation buffer for one (frames, H, W) shape and a flat list of
pre-built kernel launches; running the plan is a loop of ctypes calls on the current CUDA stream -- no allocation,
no host synchronisation, so it can also be captured into a CUDA graph.

Data layout: every activation is token-major / channels-last ([rows, C] row-major); feature maps that feed a 3x3
convolution carry a zero one-pixel border per frame ("padded", see include/vda.h).
"""
from __future__ import annotations

import ctypes as C
import os
import math

import torch

from . import lib as L
from .modules import ENCODERS

KPAD_PATCH = 640  # 3*14*14 = 588 padded to a multiple of the 64-wide K block

_graph_fallback_warned = False


def _warn_graph_fallback(exc):
    """CUDA-graph capture of a launch list failed: the plan keeps running as a loop of launches (same results, more host
    time per window).  Said once per process so that a silent slow path cannot hide in a benchmark."""
    global _graph_fallback_warned
    if not _graph_fallback_warned:
        _graph_fallback_warned = True
        import warnings
        warnings.warn(f"video_depth_anything_b200: CUDA graph capture failed ({type(exc).__name__}: {exc}); "
                      "falling back to per-kernel launches", RuntimeWarning, stacklevel=3)


def active_toggles():
    """The VDA_* environment switches that change numerics or the execution path, as set in this process (recorded in
    bench lines)."""
    return {k: v for k, v in sorted(os.environ.items()) if k.startswith("VDA_")}


def _conv3x3_pack(w: torch.Tensor) -> torch.Tensor:
    """[Cout, Cin, 3, 3] -> [Cout, 9*Cin], k = (dy*3+dx)*Cin + ci."""
    co, ci = w.shape[:2]
    return w.permute(0, 2, 3, 1).reshape(co, 9 * ci).contiguous()


def _convt_stride(cout: int) -> int:
    """Columns per (ky, kx) block of a packed transposed conv: the channel count rounded up to a multiple of 64, so
    that every 64-column epilogue chunk lies inside one block and takes the coalesced path (vits: 48 -> 64, 96 -> 128)."""
    return (cout + 63) // 64 * 64


def _convt_pack(w: torch.Tensor) -> torch.Tensor:
    """ConvTranspose2d weight [Cin, Cout, k, k] -> [(ky*k+kx)*stride + co, Cin], zero rows for co >= Cout."""
    ci, co, k, _ = w.shape
    st = _convt_stride(co)
    out = torch.zeros((k * k, st, ci), dtype=w.dtype, device=w.device)
    out[:, :co] = w.permute(2, 3, 1, 0).reshape(k * k, co, ci)
    return out.reshape(k * k * st, ci).contiguous()


def _rope_table(C_: int, max_len: int, device) -> torch.Tensor:
    """`precompute_freqs_cis(query_dim, temporal_max_len)` (motion_module/attention.py:403-408) as interleaved
    (cos, sin) fp32 pairs [max_len, C] -- the operand of vda_rope_qk."""
    freqs = 1.0 / (10000.0 ** (torch.arange(0, C_, 2)[: (C_ // 2)].float() / C_))
    ang = torch.outer(torch.arange(max_len, dtype=torch.float32), freqs)
    cis = torch.polar(torch.ones_like(ang), ang)
    return torch.view_as_real(cis).reshape(max_len, C_).contiguous().to(device)


def _geglu_interleave(t: torch.Tensor) -> torch.Tensor:
    """Rows [value(4C) ; gate(4C)] -> groups of 64 rows = 32 value rows followed by their 32 gate rows."""
    half = t.shape[0] // 2
    v, g = t[:half], t[half:]
    v = v.reshape(half // 32, 32, *t.shape[1:])
    g = g.reshape(half // 32, 32, *t.shape[1:])
    return torch.cat([v, g], dim=1).reshape(t.shape).contiguous()


class PackedWeights:
    """All weights of one model on one device in one precision (bf16 build or fp32 build)."""

    def __init__(self, model, device, bf16: bool):
        self.bf16 = bf16
        wdt = torch.bfloat16 if bf16 else torch.float32
        sd = {k: v.detach().to(device=device, dtype=torch.float32) for k, v in model.state_dict().items()}
        self.encoder = model.encoder
        enc = ENCODERS[model.encoder]
        D = enc["dim"]
        self.D, self.heads, self.depth, self.taps = D, enc["heads"], enc["depth"], enc["taps"]
        self.features = sd["head.scratch.layer1_rn.weight"].shape[0]
        self.oc = [sd[f"head.projects.{i}.weight"].shape[0] for i in range(4)]
        w = {}

        def mat(name, t):
            w[name] = t.to(wdt).contiguous()

        def vec(name, t):
            w[name] = t.float().contiguous()

        p = "pretrained."
        pw = torch.zeros(D, KPAD_PATCH, device=device)
        pw[:, :588] = sd[p + "patch_embed.proj.weight"].reshape(D, 588)
        mat("patch.w", pw)
        vec("patch.b", sd[p + "patch_embed.proj.bias"])
        vec("cls", sd[p + "cls_token"].reshape(D))
        vec("pos", sd[p + "pos_embed"].reshape(-1, D))
        for i in range(self.depth):
            b = f"{p}blocks.{i}."
            for n in ("norm1", "norm2"):
                vec(f"b{i}.{n}.w", sd[b + n + ".weight"])
                vec(f"b{i}.{n}.b", sd[b + n + ".bias"])
            for n, key in (("qkv", "attn.qkv"), ("proj", "attn.proj"), ("fc1", "mlp.fc1"), ("fc2", "mlp.fc2")):
                mat(f"b{i}.{n}.w", sd[b + key + ".weight"])
                vec(f"b{i}.{n}.b", sd[b + key + ".bias"])
            vec(f"b{i}.ls1", sd[b + "ls1.gamma"])
            vec(f"b{i}.ls2", sd[b + "ls2.gamma"])
        vec("norm.w", sd[p + "norm.weight"])
        vec("norm.b", sd[p + "norm.bias"])

        h = "head."
        for i in range(4):
            mat(f"proj{i}.w", sd[f"{h}projects.{i}.weight"].reshape(self.oc[i], D))
            vec(f"proj{i}.b", sd[f"{h}projects.{i}.bias"])
        mat("rs0.w", _convt_pack(sd[h + "resize_layers.0.weight"]))
        vec("rs0.b", sd[h + "resize_layers.0.bias"])
        mat("rs1.w", _convt_pack(sd[h + "resize_layers.1.weight"]))
        vec("rs1.b", sd[h + "resize_layers.1.bias"])
        mat("rs3.w", _conv3x3_pack(sd[h + "resize_layers.3.weight"]))
        vec("rs3.b", sd[h + "resize_layers.3.bias"])
        sc = h + "scratch."
        for i in range(1, 5):
            mat(f"rn{i}.w", _conv3x3_pack(sd[f"{sc}layer{i}_rn.weight"]))
            r = f"{sc}refinenet{i}."
            mat(f"ref{i}.out.w", sd[r + "out_conv.weight"].reshape(self.features, self.features))
            vec(f"ref{i}.out.b", sd[r + "out_conv.bias"])
            for u in (1, 2):
                for c in (1, 2):
                    mat(f"ref{i}.rcu{u}.c{c}.w", _conv3x3_pack(sd[f"{r}resConfUnit{u}.conv{c}.weight"]))
                    vec(f"ref{i}.rcu{u}.c{c}.b", sd[f"{r}resConfUnit{u}.conv{c}.bias"])
        mat("oc1.w", _conv3x3_pack(sd[sc + "output_conv1.weight"]))
        vec("oc1.b", sd[sc + "output_conv1.bias"])
        # the fp32 island (dpt_temporal.py:95-97): weights stay fp32 (tcgen05 kind::tf32 in the bf16 build)
        w2 = _conv3x3_pack(sd[sc + "output_conv2.0.weight"]).float()
        w["oc2.w32"] = w2.contiguous()
        w["oc2.w16"] = w2.to(torch.float16).contiguous()  # operands of vda_output_head_f16 (round to nearest, 10-bit mantissa like TF32)
        # fp16 operands only when the WEIGHTS fit fp16: no overflow, and the absolute rounding step of fp16's subnormal
        # range (2^-25) negligible against the largest weight (the activations are range-checked on the device, per launch)
        wmax = float(w2.abs().max().item())
        self.oc2_fp16_safe = bool(torch.isfinite(w2).all().item()) and 1e-3 <= wmax <= 6.0e4
        vec("oc2.b", sd[sc + "output_conv2.0.bias"])
        vec("oc2.hw", sd[sc + "output_conv2.2.weight"].reshape(32))
        self.head_b = float(sd[sc + "output_conv2.2.bias"].reshape(()).item())
        # pe='rope' models carry no pos_encoder buffers (motion_module.py:226-240); the table length is num_frames
        self.rope = getattr(model, "pe", "ape") == "rope"
        pe_key = f"{h}motion_modules.0.temporal_transformer.transformer_blocks.0.attention_blocks.0.pos_encoder.pe"
        self.pe_len = int(model.num_frames) if self.rope else int(sd[pe_key].shape[1])
        for m in range(4):
            t = f"{h}motion_modules.{m}.temporal_transformer."
            C_m = int(sd[t + "norm.weight"].shape[0])
            vec(f"mm{m}.gn.w", sd[t + "norm.weight"])
            vec(f"mm{m}.gn.b", sd[t + "norm.bias"])
            mat(f"mm{m}.in.w", sd[t + "proj_in.weight"])
            vec(f"mm{m}.in.b", sd[t + "proj_in.bias"])
            tb = t + "transformer_blocks.0."
            for a in (0, 1):
                ab = f"{tb}attention_blocks.{a}."
                mat(f"mm{m}.a{a}.qkv.w", torch.cat([sd[ab + "to_q.weight"], sd[ab + "to_k.weight"], sd[ab + "to_v.weight"]], 0))
                mat(f"mm{m}.a{a}.out.w", sd[ab + "to_out.0.weight"])
                vec(f"mm{m}.a{a}.out.b", sd[ab + "to_out.0.bias"])
                if self.rope:
                    w[f"mm{m}.a{a}.rope"] = _rope_table(C_m, self.pe_len, device)
                else:
                    vec(f"mm{m}.a{a}.pe", sd[ab + "pos_encoder.pe"][0])
                vec(f"mm{m}.a{a}.ln.w", sd[f"{tb}norms.{a}.weight"])
                vec(f"mm{m}.a{a}.ln.b", sd[f"{tb}norms.{a}.bias"])
            mat(f"mm{m}.ff1.w", _geglu_interleave(sd[tb + "ff.net.0.proj.weight"]))
            vec(f"mm{m}.ff1.b", _geglu_interleave(sd[tb + "ff.net.0.proj.bias"]))
            mat(f"mm{m}.ff2.w", sd[tb + "ff.net.2.weight"])
            vec(f"mm{m}.ff2.b", sd[tb + "ff.net.2.bias"])
            vec(f"mm{m}.ffln.w", sd[tb + "ff_norm.weight"])
            vec(f"mm{m}.ffln.b", sd[tb + "ff_norm.bias"])
            mat(f"mm{m}.po.w", sd[t + "proj_out.weight"])
            vec(f"mm{m}.po.b", sd[t + "proj_out.bias"])
        self.w = w


class ForwardPlan:
    """Buffers + launch list for `VideoDepthAnything.forward` on x[frames=B*T, 3, H, W] (video_depth.py:58-65)."""

    # fusion blocks: VDA_FUSION_CONV_FIRST=1 runs the 1x1 out_conv before the interpolation (see _fusion).  Off by default:
    # it saves 0.5 of 56.8 ms per vitl clip but rounds in a different place than the reference, so the bf16 build's noise
    # is no longer correlated with the reference's (vitl T3 28x42: max deviation from the autocast reference 3.5e-2 ->
    # 4.6e-2 against a gate of 4.0e-2, same accuracy against fp32).
    conv_first = os.environ.get("VDA_FUSION_CONV_FIRST", "0") == "1"

    def __init__(self, pw: PackedWeights, batch: int, T: int, H: int, W: int, device, skip_tmp_block=False, use_tc=None,
                 keep=(), attn_tc=None, fused_tail=True):
        assert H % 14 == 0 and W % 14 == 0, f"Input image size {H}x{W} is not a multiple of patch size 14"  # patch_embed.py:73-74
        if T > pw.pe_len:
            raise ValueError(f"video length {T} exceeds the positional-encoding length {pw.pe_len} (num_frames)")
        self.pw, self.device = pw, device
        self.bf16 = pw.bf16
        self.adt = torch.bfloat16 if pw.bf16 else torch.float32
        self.dcode = L.BF16 if pw.bf16 else L.F32
        self.use_tc = bool(pw.bf16) if use_tc is None else bool(use_tc)
        self.attn_tc = self.use_tc if attn_tc is None else bool(attn_tc)
        self.fused_tail = bool(fused_tail)
        self.batch, self.T, self.H, self.W = batch, T, H, W
        self.BT = batch * T
        self.ph, self.pw_ = H // 14, W // 14
        self.ops = []      # (callable, args) in launch order
        self.marks = {}    # stage name -> index of its first op (stage-by-stage parity tests run `run_range` between marks)
        self.bufs = {}     # name -> tensor (kept alive; `keep` names are exposed for tap-level tests)
        self.keep = set(keep)
        self.n_launches = 0
        self._ws = None
        self._ws_need = 0
        self._lib = L.load()
        self.use_graph = os.environ.get("VDA_NO_GRAPH", "0") != "1"
        self._graph = None
        self._eager_runs = 0
        self._prefix = ""   # name prefix of the buffers being built (stages that repeat a builder need distinct names)
        self._alias = {}    # name -> caller-provided tensor to use instead of allocating (streaming mode's feature store)
        self._build(skip_tmp_block)
        if not self.use_tc:
            self._ws = torch.empty(max(self._ws_need, 4), dtype=torch.uint8, device=device)
            for fn, args in self._all_ops():
                if fn is self._lib.vda_gemm:
                    args[0].workspace = self._ws.data_ptr()
                    args[0].workspace_bytes = self._ws.numel()

    def _all_ops(self):
        return self.ops

    # ------------------------------------------------------------------ buffer helpers
    def buf(self, name, rows, cols, dtype=None, zero=False):
        dtype = dtype or self.adt
        if name in self._alias:
            t = self._alias.pop(name)
            assert tuple(t.shape) == (rows, cols) and t.dtype == dtype and t.is_contiguous(), (name, t.shape, rows, cols)
        else:
            t = (torch.zeros if zero else torch.empty)((rows, cols), dtype=dtype, device=self.device)
        assert self._prefix + name not in self.bufs, self._prefix + name  # a dropped reference would free a live buffer
        self.bufs[self._prefix + name] = t
        return t

    def padded(self, name, h, w, C):
        """Zero-bordered channels-last map [BT*(h+2)*(w+2), C]; borders are never written afterwards."""
        return self.buf(name, self.BT * (h + 2) * (w + 2), C, zero=True)

    # ------------------------------------------------------------------ op builders
    def gemm(self, A, B, M, N, K, *, rowsA=None, lda=None, taps=1, ksplit=1, geo=None, m_padded=0, bias=None,
             act=L.ACT_NONE, out=None, ldo=None, out_map=L.MAP_SAME, out_relu=None, res=None, res2=None, ldres=None,
             res_map=L.MAP_SAME, resid=None, gamma=None, pos=None, resid_map=L.MAP_SAME, resid_assign=0, pos_rows=0,
             ct_k=0, ct_cout=0, head_w=None, head_out=None, head_b=0.0, use_tc=None, b_dtype=None):
        d = L.GemmDesc()
        d.A, d.B = A.data_ptr(), B.data_ptr()
        d.lda = lda if lda is not None else A.shape[1]
        d.ldb = B.shape[1]
        d.rowsA = rowsA if rowsA is not None else A.shape[0]
        d.a_dtype = L.dt(A)
        d.b_dtype = L.dt(B) if b_dtype is None else b_dtype
        d.M, d.N, d.K, d.taps, d.ksplit = M, N, K, taps, ksplit
        if geo is not None:
            d.gb, d.gh, d.gw = geo
        d.m_padded = m_padded
        d.bias = L._ptr(bias)
        d.act = act
        d.round_bf16 = 1 if self.bf16 else 0
        d.out_dtype = self.dcode
        d.out = L._ptr(out)
        d.out_relu = L._ptr(out_relu)
        ncols = N // 2 if act == L.ACT_GEGLU else (ct_cout if ct_k else N)
        d.ldo = ldo if ldo is not None else (out.shape[1] if out is not None else ncols)
        d.out_map = out_map
        d.ct_k, d.ct_cout = ct_k, ct_cout
        d.ct_stride = _convt_stride(ct_cout) if ct_k else 0
        d.res, d.res2 = L._ptr(res), L._ptr(res2)
        d.ldres = ldres if ldres is not None else (res.shape[1] if res is not None else d.ldo)
        d.res_map = res_map
        d.resid, d.gamma, d.pos = L._ptr(resid), L._ptr(gamma), L._ptr(pos)
        d.ldresid = resid.shape[1] if resid is not None else 0
        d.resid_map, d.resid_assign, d.pos_rows = resid_map, resid_assign, pos_rows
        d.head_w, d.head_out, d.head_b = L._ptr(head_w), L._ptr(head_out), head_b
        d.use_tc = int(self.use_tc if use_tc is None else use_tc)
        if not d.use_tc:
            self._ws_need = max(self._ws_need, M * N * 4)
        self.ops.append((self._lib.vda_gemm, [d]))
        self.n_launches += 1 if d.use_tc else 2
        return d

    def mark(self, name):
        self.marks[self._prefix + name] = len(self.ops)

    def run_range(self, first, last):
        """Launch ops[first:last] on the current stream (tests: one stage of the network on caller-written inputs)."""
        keep = self.ops
        self.ops = keep[first:last]
        try:
            self._launch_ops()
        finally:
            self.ops = keep

    def op(self, fn, *args, launches=1):
        self.ops.append((fn, list(args)))
        self.n_launches += launches

    def layernorm(self, x, y, w, b, eps, rows, D, group_rows=0, skip_first=0, pe=None, pe_group=0, pe_T=0):
        self.op(self._lib.vda_layernorm, x.data_ptr(), L.dt(x), y.data_ptr(), L.dt(y), w.data_ptr(), b.data_ptr(), eps,
                rows, D, group_rows, skip_first, L._ptr(pe), pe_group, pe_T)

    def conv3x3(self, xpad, wname, h, w, Cin, Cout, *, bias=None, **kw):
        """3x3 s1 p1 conv over a padded map as a 9-tap GEMM (M runs over the padded grid)."""
        M = self.BT * (h + 2) * (w + 2)
        return self.gemm(xpad, self.pw.w[wname], M, Cout, Cin, taps=9, geo=(self.BT, h, w), m_padded=1, bias=bias, **kw)

    # ------------------------------------------------------------------ the network
    def _build(self, skip_tmp_block):
        self._build_head(self._build_encoder(), skip_tmp_block)

    def _build_encoder(self):
        """Patch embed + transformer blocks + the final norm on the four taps (dinov2.py:297-321).  Returns the taps."""
        pw, W_ = self.pw, self.pw.w
        BT, ph, pwid, D = self.BT, self.ph, self.pw_, pw.D
        P = ph * pwid
        ntok = P + 1
        M = BT * ntok
        lib = self._lib
        self.x_in = torch.empty((BT, 3, self.H, self.W), dtype=torch.float32, device=self.device)
        # The patch embed reads frame f at src_ptrs[f] (a device array, read when the kernel runs).  By default these are
        # the frames of x_in; the sliding-window driver points them at its pool of once-pre-processed frames instead, per
        # window, so a window's inputs (video_depth.py:358-364) are gathered without a copy, also under graph replay.
        fbytes = 3 * self.H * self.W * 4
        self.src_ptrs_default = torch.tensor([self.x_in.data_ptr() + f * fbytes for f in range(BT)], dtype=torch.int64).to(self.device)
        self.src_ptrs = self.src_ptrs_default.clone()

        # ---- position embedding for this resolution (dinov2.py:179-210): identity at 518x518, else bicubic
        if P == 37 * 37 and self.H == self.W:
            pos = W_["pos"]
        else:
            pos = torch.empty((ntok, D), dtype=torch.float32, device=self.device)
            check = L.check
            check(lib.vda_pos_embed_bicubic(W_["pos"].data_ptr(), pos.data_ptr(), D, 37, ph, pwid,
                                            (ph + 0.1) / 37.0, (pwid + 0.1) / 37.0, L.stream_ptr()), "vda_pos_embed_bicubic")
        self.bufs["pos"] = pos

        # ---- patch embed (patch_embed.py:76-78) + cls + pos (dinov2.py:212-221)
        a_pe = self.buf("im2col", BT * P, KPAD_PATCH)
        resid = self.buf("resid", M, D, torch.float32)
        self.op(lib.vda_im2col_patch_frames, self.src_ptrs.data_ptr(), a_pe.data_ptr(), self.dcode, BT, self.H, self.W, KPAD_PATCH)
        self.op(lib.vda_cls_rows, resid.data_ptr(), W_["cls"].data_ptr(), pos.data_ptr(), BT, ntok, D)
        self.gemm(a_pe, W_["patch.w"], BT * P, D, KPAD_PATCH, geo=(BT, ph, pwid), bias=W_["patch.b"], resid=resid, pos=pos,
                  resid_map=L.MAP_TOKENS, resid_assign=1, pos_rows=ntok)
        if "tokens0" in self.keep:
            self.bufs["tokens0"] = resid

        # ---- transformer blocks (block.py:83-87,105-106)
        y = self.buf("y", M, D)
        qkv = self.buf("qkv", M, 3 * D)
        att = self.buf("att", M, D)
        hid = self.buf("hid", M, 4 * D)
        taps = []
        for i in range(pw.depth):
            self.mark(f"blk{i}")
            self.layernorm(resid, y, W_[f"b{i}.norm1.w"], W_[f"b{i}.norm1.b"], 1e-6, M, D)
            self.gemm(y, W_[f"b{i}.qkv.w"], M, 3 * D, D, bias=W_[f"b{i}.qkv.b"], out=qkv)
            self.op(lib.vda_attention, qkv.data_ptr(), att.data_ptr(), self.dcode, BT, ntok, pw.heads, int(self.bf16),
                    int(self.attn_tc))
            self.gemm(att, W_[f"b{i}.proj.w"], M, D, D, bias=W_[f"b{i}.proj.b"], resid=resid, gamma=W_[f"b{i}.ls1"])
            self.layernorm(resid, y, W_[f"b{i}.norm2.w"], W_[f"b{i}.norm2.b"], 1e-6, M, D)
            self.gemm(y, W_[f"b{i}.fc1.w"], M, 4 * D, D, bias=W_[f"b{i}.fc1.b"], act=L.ACT_GELU, out=hid)
            self.gemm(hid, W_[f"b{i}.fc2.w"], M, D, 4 * D, bias=W_[f"b{i}.fc2.b"], resid=resid, gamma=W_[f"b{i}.ls2"])
            if i in pw.taps:
                self.mark(f"tapnorm{len(taps)}")
                t = self.buf(f"tap{len(taps)}", BT * P, D)
                # final norm on the tap, class token dropped (dinov2.py:309-312)
                self.layernorm(resid, t, W_["norm.w"], W_["norm.b"], 1e-6, M, D, group_rows=ntok, skip_first=1)
                taps.append(t)
        self.mark("enc_end")
        return taps

    # ------------------------------------------------------------------ motion module (motion_module.py:108-133,173-186)
    def _motion_module(self, m, x, hw_geo, C, out, out_map):
        """x: dense tokens [BT*h*w, C] (also the residual).  out: dense or padded buffer receiving proj_out(h) + x."""
        att, hbuf = self._mm_front(m, x, hw_geo, C)
        self._mm_back(m, att, hbuf, x, hw_geo, C, out, out_map)

    def _mm_front(self, m, x, hw_geo, C):
        """The temporally coupled part of motion module m: GroupNorm, proj_in, the first attention block and the second
        one up to its attention output (motion_module.py:141-164,187-200).  Returns (att, h): the second block's
        attention output and the residual stream it will be projected onto; everything after is per frame (_mm_back)."""
        pw, W_, lib = self.pw, self.pw.w, self._lib
        h_, w_ = hw_geo
        hw = h_ * w_
        M = self.BT * hw
        n = f"mm{m}"
        g = self.buf(n + ".g", M, C)
        hbuf = self.buf(n + ".h", M, C)
        nbuf = self.buf(n + ".n", M, C)
        qkv = self.buf(n + ".qkv", M, 3 * C)
        att = self.buf(n + ".att", M, C)
        stats = torch.empty((self.BT * 32 * 2 * 34,), dtype=torch.float64, device=self.device)
        self.bufs[self._prefix + n + ".stats"] = stats
        self.op(lib.vda_groupnorm_tokens, x.data_ptr(), self.dcode, g.data_ptr(), self.dcode, W_[n + ".gn.w"].data_ptr(),
                W_[n + ".gn.b"].data_ptr(), 1e-6, self.BT, hw, C, 32, stats.data_ptr(), launches=3)
        self.gemm(g, W_[n + ".in.w"], M, C, C, bias=W_[n + ".in.b"], out=hbuf)
        for a in (0, 1):
            if pw.rope:
                self.layernorm(hbuf, nbuf, W_[f"{n}.a{a}.ln.w"], W_[f"{n}.a{a}.ln.b"], 1e-5, M, C)
            else:
                self.layernorm(hbuf, nbuf, W_[f"{n}.a{a}.ln.w"], W_[f"{n}.a{a}.ln.b"], 1e-5, M, C, pe=W_[f"{n}.a{a}.pe"],
                               pe_group=hw, pe_T=self.T)
            self.gemm(nbuf, W_[f"{n}.a{a}.qkv.w"], M, 3 * C, C, out=qkv)
            if pw.rope:  # rotate q and k in place (motion_module.py:290-293)
                self.op(lib.vda_rope_qk, qkv.data_ptr(), self.dcode, W_[f"{n}.a{a}.rope"].data_ptr(), M, hw, self.T, C,
                        int(self.bf16))
            self.op(lib.vda_temporal_attention, qkv.data_ptr(), att.data_ptr(), self.dcode, self.batch, self.T, hw, C, 8,
                    int(self.bf16))
            if a == 0:
                self.gemm(att, W_[f"{n}.a{a}.out.w"], M, C, C, bias=W_[f"{n}.a{a}.out.b"], out=hbuf, res=hbuf)
        return att, hbuf

    def _mm_back(self, m, att, hbuf, x, hw_geo, C, out, out_map):
        """The per-frame rest of motion module m over self.BT frames: the second attention block's output projection,
        the feed-forward, proj_out and the module's residual (motion_module.py:165-172,201-210).  hbuf is updated in place."""
        W_ = self.pw.w
        h_, w_ = hw_geo
        M = self.BT * h_ * w_
        n = f"mm{m}"
        nbuf = self.bufs.get(self._prefix + n + ".n")  # the front's LayerNorm buffer when both halves run in one stage
        if nbuf is None or tuple(nbuf.shape) != (M, C):
            nbuf = self.buf(n + ".n2", M, C)
        ff = self.buf(n + ".ff", M, 4 * C)
        self.gemm(att, W_[f"{n}.a1.out.w"], M, C, C, bias=W_[f"{n}.a1.out.b"], out=hbuf, res=hbuf)
        self.layernorm(hbuf, nbuf, W_[n + ".ffln.w"], W_[n + ".ffln.b"], 1e-5, M, C)
        self.gemm(nbuf, W_[n + ".ff1.w"], M, 8 * C, C, bias=W_[n + ".ff1.b"], act=L.ACT_GEGLU, out=ff)
        self.gemm(ff, W_[n + ".ff2.w"], M, C, 4 * C, bias=W_[n + ".ff2.b"], out=hbuf, res=hbuf)
        self.gemm(hbuf, W_[n + ".po.w"], M, C, C, geo=(self.BT, h_, w_), bias=W_[n + ".po.b"], out=out, out_map=out_map, res=x,
                  res_map=L.MAP_SAME)
        if n in self.keep:
            self.bufs[n] = out

    # ------------------------------------------------------------------ fusion blocks (util/blocks.py:68-91,135-162)
    def _rcu(self, name, x, x_relu, h, w, F, out, out_relu=None, res2=None):
        """out = conv2(relu(conv1(relu(x)))) + x (+ res2); all maps padded at (h, w)."""
        W_ = self.pw.w
        t = self.padded(name + ".t", h, w, F)
        self.conv3x3(x_relu, name + ".c1.w", h, w, F, F, bias=W_[name + ".c1.b"], act=L.ACT_RELU, out=t)
        self.conv3x3(t, name + ".c2.w", h, w, F, F, bias=W_[name + ".c2.b"], out=out, out_relu=out_relu, res=x, res2=res2)

    def _fusion(self, i, x0, x0_relu, x1, x1_relu, h, w, oh, ow, out, out_map):
        """refinenet{i}: x0 (+ RCU1(x1)) -> RCU2 -> bilinear to (oh, ow) -> out_conv.  x0/x1 padded at (h, w)."""
        W_, F = self.pw.w, self.pw.features
        n = f"ref{i}"
        if x1 is not None:
            s = self.padded(n + ".s", h, w, F)
            s_relu = self.padded(n + ".s_relu", h, w, F)
            self._rcu(n + ".rcu1", x1, x1_relu, h, w, F, s, out_relu=s_relu, res2=x0)
        else:
            s, s_relu = x0, x0_relu
        r = self.padded(n + ".r", h, w, F)
        self._rcu(n + ".rcu2", s, s_relu, h, w, F, r)
        if self.conv_first:
            # the 1x1 out_conv and the interpolation are both linear and the bilinear weights sum to one (bias included),
            # so they commute: the conv runs on the (h, w) map -- a quarter of the rows when upsampling 2x -- and the
            # interpolation writes the block's output.  The conv's rows on the zero ring hold the bias and are never read
            # (align_corners interpolation stays inside the map).
            c = self.buf(n + ".c", self.BT * (h + 2) * (w + 2), F)
            self.gemm(r, W_[n + ".out.w"], self.BT * (h + 2) * (w + 2), F, F, bias=W_[n + ".out.b"], out=c)
            assert out_map in (L.MAP_PADDED, L.MAP_SAME)
            self.op(self._lib.vda_bilinear, c.data_ptr(), self.dcode, out.data_ptr(), self.dcode, self.BT, h, w, 1, oh, ow,
                    1 if out_map == L.MAP_PADDED else 0, F)
            return
        u = self.buf(n + ".u", self.BT * oh * ow, F)
        self.op(self._lib.vda_bilinear, r.data_ptr(), self.dcode, u.data_ptr(), self.dcode, self.BT, h, w, 1, oh, ow, 0, F)
        self.gemm(u, W_[n + ".out.w"], self.BT * oh * ow, F, F, geo=(self.BT, oh, ow), bias=W_[n + ".out.b"], out=out,
                  out_map=out_map)

    # ------------------------------------------------------------------ DPTHeadTemporal.forward (dpt_temporal.py:53-99)
    # The head is built in three stages so that the streaming mode (dpt_temporal.py:132-260) can run them on different
    # frame counts: reassemble (the new frame), middle (the T-frame temporal context), tail (the frames to predict).
    def _reassemble(self, taps):
        """projects (1x1) + resize layers (dpt.py:50-90; dpt_temporal.py:55-69).  Returns layer_1, layer_2 (padded) and
        layer_3, layer_4 (dense)."""
        pw, W_ = self.pw, self.pw.w
        BT, ph, pwid, D, oc = self.BT, self.ph, self.pw_, pw.D, pw.oc
        P = ph * pwid
        geo3 = (BT, ph, pwid)
        h1, w1, h2, w2, h3, w3 = 4 * ph, 4 * pwid, 2 * ph, 2 * pwid, ph, pwid
        h4, w4 = (ph + 1) // 2, (pwid + 1) // 2
        t0 = self.buf("t0", BT * P, oc[0])
        self.gemm(taps[0], W_["proj0.w"], BT * P, oc[0], D, bias=W_["proj0.b"], out=t0)
        l1 = self.padded("layer_1", h1, w1, oc[0])
        self.gemm(t0, W_["rs0.w"], BT * P, 16 * _convt_stride(oc[0]), oc[0], geo=geo3, bias=W_["rs0.b"], out=l1, out_map=L.MAP_CONVT, ct_k=4,
                  ct_cout=oc[0])
        t1 = self.buf("t1", BT * P, oc[1])
        self.gemm(taps[1], W_["proj1.w"], BT * P, oc[1], D, bias=W_["proj1.b"], out=t1)
        l2 = self.padded("layer_2", h2, w2, oc[1])
        self.gemm(t1, W_["rs1.w"], BT * P, 4 * _convt_stride(oc[1]), oc[1], geo=geo3, bias=W_["rs1.b"], out=l2, out_map=L.MAP_CONVT, ct_k=2,
                  ct_cout=oc[1])
        l3 = self.buf("layer_3", BT * P, oc[2])
        self.gemm(taps[2], W_["proj2.w"], BT * P, oc[2], D, bias=W_["proj2.b"], out=l3)
        # resize_layers[3] is a stride-2 3x3 conv (dpt.py:83-88): projects[3] scatters its output into the four parity
        # planes of the padded map (VDA_MAP_S2D), the conv then runs over ONE plane grid with per-tap plane offsets
        # (m_padded = 2) and computes only the outputs that are kept -- a quarter of the rows of the padded-grid form
        hp, wp = h4 + 1, w4 + 1
        t3 = self.buf("t3", 4 * BT * hp * wp, oc[3], zero=True)
        self.gemm(taps[3], W_["proj3.w"], BT * P, oc[3], D, geo=geo3, bias=W_["proj3.b"], out=t3, out_map=L.MAP_S2D)
        l4 = self.buf("layer_4", BT * h4 * w4, oc[3])
        self.gemm(t3, W_["rs3.w"], BT * hp * wp, oc[3], oc[3], taps=9, geo=(BT, hp, wp), m_padded=2, bias=W_["rs3.b"], out=l4,
                  out_map=L.MAP_CROP)
        return l1, l2, l3, l4

    def _rn(self, i, src, h, w):
        """scratch.layer{i}_rn (3x3, no bias; blocks.py:20-33): raw + relu copies, both padded."""
        F, oc = self.pw.features, self.pw.oc
        raw = self.padded(f"l{i}r", h, w, F)
        relu = self.padded(f"l{i}r_relu", h, w, F)
        self.conv3x3(src, f"rn{i}.w", h, w, oc[i - 1], F, out=raw, out_relu=relu)
        return raw, relu

    def _head_middle(self, l3, l4, skip_tmp_block, split_mm3=False):
        """Motion modules 0/1, layer3/4_rn, refinenet4 (+ motion module 2), refinenet3 + motion module 3
        (dpt_temporal.py:75-76,80-87) over self.BT = batch * T frames.  Returns path_3 (padded at the layer_2 size), or with
        split_mm3 the inputs of motion module 3's per-frame part (att, h, x; dense at the layer_2 size) -- the streaming mode
        finishes that module only for the frames it predicts."""
        pw = self.pw
        ph, pwid, F, oc = self.ph, self.pw_, pw.features, pw.oc
        BT = self.BT
        h2, w2, h3, w3 = 2 * ph, 2 * pwid, ph, pwid
        h4, w4 = (ph + 1) // 2, (pwid + 1) // 2
        self.mark("mm0")
        l3m = self.padded("mm0", h3, w3, oc[2])
        self._motion_module(0, l3, (h3, w3), oc[2], l3m, L.MAP_PADDED)
        self.mark("mm1")
        l4m = self.padded("mm1", h4, w4, oc[3])
        self._motion_module(1, l4, (h4, w4), oc[3], l4m, L.MAP_PADDED)
        self.mark("rn34_f4")
        rn3, rn3_relu = self._rn(3, l3m, h3, w3)
        rn4, rn4_relu = self._rn(4, l4m, h4, w4)
        if skip_tmp_block:
            p4 = self.padded("path_4", h3, w3, F)
            self._fusion(4, rn4, rn4_relu, None, None, h4, w4, h3, w3, p4, L.MAP_PADDED)
        else:
            p4d = self.buf("path_4d", BT * h3 * w3, F)
            self._fusion(4, rn4, rn4_relu, None, None, h4, w4, h3, w3, p4d, L.MAP_SAME)
            self.mark("mm2")
            p4 = self.padded("path_4", h3, w3, F)
            self._motion_module(2, p4d, (h3, w3), F, p4, L.MAP_PADDED)
        self.mark("f3")
        p3d = self.buf("path_3d", BT * h2 * w2, F)
        self._fusion(3, p4, None, rn3, rn3_relu, h3, w3, h2, w2, p3d, L.MAP_SAME)
        self.mark("mm3")
        if split_mm3:
            att, hbuf = self._mm_front(3, p3d, (h2, w2), F)
            return att, hbuf, p3d
        p3 = self.padded("path_3", h2, w2, F)
        self._motion_module(3, p3d, (h2, w2), F, p3, L.MAP_PADDED)
        return p3

    def _head_tail(self, p3, l1, l2):
        """layer1/2_rn, refinenet2, refinenet1, output_conv1, upsample, the fp32 island output_conv2
        (dpt_temporal.py:78-79,88-99) over self.BT frames.  Returns depth [BT, H, W] fp32."""
        pw, W_, lib = self.pw, self.pw.w, self._lib
        BT, ph, pwid, F = self.BT, self.ph, self.pw_, pw.features
        h1, w1, h2, w2 = 4 * ph, 4 * pwid, 2 * ph, 2 * pwid
        self.mark("rn12_f2")
        rn1, rn1_relu = self._rn(1, l1, h1, w1)
        rn2, rn2_relu = self._rn(2, l2, h2, w2)
        p2 = self.padded("path_2", h1, w1, F)
        self._fusion(2, p3, None, rn2, rn2_relu, h2, w2, h1, w1, p2, L.MAP_PADDED)
        self.mark("f1")
        h0, w0 = 2 * h1, 2 * w1
        p1 = self.padded("path_1", h0, w0, F)
        self._fusion(1, p2, None, rn1, rn1_relu, h1, w1, h0, w0, p1, L.MAP_PADDED)
        self.mark("out")
        # output_conv1 -> bilinear to (14ph, 14pw) -> fp32 island output_conv2 (dpt_temporal.py:91-99)
        o1 = self.buf("oc1", BT * h0 * w0, F // 2)
        self.conv3x3(p1, "oc1.w", h0, w0, F, F // 2, bias=W_["oc1.b"], out=o1, out_map=L.MAP_DENSE)
        # upsample_bilinear2d is an fp32 op under CUDA autocast: the upsampled map is fp32 and un-rounded in both builds, and
        # so is everything after it (ori_type is fp32, so `out.to(ori_type)` at dpt_temporal.py:99 is a no-op)
        depth = torch.empty((BT, self.H, self.W), dtype=torch.float32, device=self.device)
        self.bufs[self._prefix + "depth"] = depth
        if self.use_tc and self.fused_tail and (F // 2) % 32 == 0:
            # bf16 build: interpolation, TF32 tensor-core conv and the ReLU -> 1x1 -> ReLU tail in one kernel
            # fp16 operands (same 10 mantissa bits as TF32, rounded to nearest, half the MMAs) when the weights fit fp16,
            # behind a device-side range guard on the activations that falls back to the TF32 kernel: nothing saturates
            # silently (dpt_temporal.py:95-97 runs this conv in fp32 / TF32)
            half = (F // 2) % 64 == 0 and pw.oc2_fp16_safe and os.environ.get("VDA_HEAD_TF32", "0") != "1"
            if half:
                self.head_flag = torch.zeros(1, dtype=torch.int32, device=self.device)
                self.bufs[self._prefix + "head_flag"] = self.head_flag
                self.op(lib.vda_output_head_guarded, o1.data_ptr(), W_["oc2.w16"].data_ptr(), W_["oc2.w32"].data_ptr(),
                        W_["oc2.b"].data_ptr(), W_["oc2.hw"].data_ptr(), float(pw.head_b), depth.data_ptr(), BT, h0, w0, F // 2,
                        self.H, self.W, self.head_flag.data_ptr(), launches=3)
            else:
                self.op(lib.vda_output_head, o1.data_ptr(), W_["oc2.w32"].data_ptr(), W_["oc2.b"].data_ptr(), W_["oc2.hw"].data_ptr(),
                        float(pw.head_b), depth.data_ptr(), BT, h0, w0, F // 2, self.H, self.W)
        else:
            up = self.buf("up", BT * (self.H + 2) * (self.W + 2), F // 2, torch.float32, zero=True)
            self.op(lib.vda_bilinear, o1.data_ptr(), self.dcode, up.data_ptr(), L.F32, BT, h0, w0, 0, self.H, self.W, 1, F // 2)
            Mup = BT * (self.H + 2) * (self.W + 2)
            # fp32 operands: SIMT fp32 FMA in the fp32 build, tcgen05 kind::tf32 in the bf16 build (cuDNN's default for fp32 convs)
            self.gemm(up, W_["oc2.w32"], Mup, 32, F // 2, taps=9, geo=(BT, self.H, self.W), m_padded=1, bias=W_["oc2.b"],
                      out_map=L.MAP_DENSE, head_w=W_["oc2.hw"], head_out=depth, head_b=pw.head_b)
        return depth

    def _build_head(self, taps, skip_tmp_block):
        self.mark("reassemble")
        l1, l2, l3, l4 = self._reassemble(taps)
        p3 = self._head_middle(l3, l4, skip_tmp_block)
        self.depth = self._head_tail(p3, l1, l2)
        self.mark("end")

    # ------------------------------------------------------------------ execution
    def run(self, x: torch.Tensor) -> torch.Tensor:
        """x: [BT,3,H,W] fp32 on this plan's device.  Returns the plan-owned depth buffer [BT,H,W]."""
        self.x_in.copy_(x.reshape(self.x_in.shape), non_blocking=True)
        self.launch()
        return self.depth

    def launch(self):
        """Enqueue the whole forward on the current stream.  After the first (eager) run the launch list is captured
        into a CUDA graph and replayed: the list is static (fixed buffers, tensor maps passed by value), and for the
        small encoders the ~270 ctypes calls per window otherwise cost as much host time as the GPU needs."""
        if self._graph is not None:
            self._graph.replay()
            return
        self._launch_ops()
        self._eager_runs += 1
        if self.use_graph and self._eager_runs == 2 and not torch.cuda.is_current_stream_capturing():
            g = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(g):
                    self._launch_ops()
                self._graph = g
            except Exception as e:  # capture is an optimisation: stay on the launch loop if the runtime refuses
                self.use_graph = False
                _warn_graph_fallback(e)

    def _launch_ops(self):
        st = L.stream_ptr()
        lib = self._lib
        gemm = lib.vda_gemm
        for fn, args in self.ops:
            if fn is gemm:
                rc = fn(C.byref(args[0]), st)
            else:
                rc = fn(*args, st)
            if rc != 0:
                raise L.VdaError(f"{fn.__name__} failed (rc={rc}): {lib.vda_last_error().decode(errors='replace')}")


class HeadPlan(ForwardPlan):
    """`DPTHeadTemporal.forward` alone (dpt_temporal.py:53-99): the four encoder taps [frames*ph*pw, D] are inputs
    (`self.taps`, written by the caller), the output is `self.depth`.  BASELINE.json config 5 (isolated head / motion
    module timing at 16 / 32 / 64 frames) runs on this plan."""

    def _build(self, skip_tmp_block):
        P = self.ph * self.pw_
        self.taps = [self.buf(f"tap{i}", self.BT * P, self.pw.D) for i in range(4)]
        self._build_head(self.taps, skip_tmp_block)


class StreamPlan(ForwardPlan):
    """Launch lists of the streaming single-image mode (reference `forward_single_image` / `foward_single_image`,
    video_depth.py:67-91, dpt_temporal.py:132-260) for one network size.

    The reference keeps the reassembled features (layer_1..4) of past frames in four host-managed tensors, re-gathers
    T-1 of them by fancy indexing every frame and shifts the whole store with a copy once it is full
    (video_depth.py:244-258,281-296).  Here the store is four device rings of `rows` slots + one incoming slot; the
    encoder stage writes the new frame's features straight into the incoming slot, the context is gathered by
    `vda_copy_frames` through a device index array (so every stage is a static launch list, replayable as a CUDA graph),
    and "shifting" the store is an index-table update on the host.

    Stages:  "encode" (`enc_batch` frames: encoder + projects/resize layers)  ->  "middle" (T context frames: motion modules
    0-3, layer3/4_rn, refinenet4/3)  ->  "tail{n}" (the n frames to predict: layer1/2_rn, refinenet2/1, output convs)
    -> "commit" (incoming slot -> its ring slot).
    """

    IDX_CTX, IDX_PRED, IDX_P3, IDX_COMMIT, IDX_SRC = range(5)

    def __init__(self, pw: PackedWeights, T: int, H: int, W: int, device, rows: int, npreds=(1,), skip_tmp_block=False,
                 use_tc=None, attn_tc=None, enc_batch=1):
        self.rows = int(rows)
        # The encoder has no temporal coupling, so when the frames are all in memory (they are: the API takes the whole
        # video) the encode stage runs on `enc_batch` consecutive frames at once into as many incoming slots -- the same
        # per-frame arithmetic, on GEMMs that fill the machine (one 518x518 frame is 44 tiles for 148 SMs).
        self.enc_batch = max(1, int(enc_batch))
        self.npreds = tuple(sorted(set(int(n) for n in npreds)))
        self.stages = {}
        super().__init__(pw, 1, T, H, W, device, skip_tmp_block, use_tc=use_tc, attn_tc=attn_tc)

    def _all_ops(self):
        return [op for name, st in self.stages.items() if name != "reassemble" for op in st["ops"]]   # a slice of "encode"

    def _end_stage(self, name):
        self.stages[name] = {"ops": self.ops, "graph": None, "runs": 0}
        self.ops = []

    def _build(self, skip_tmp_block):
        pw, lib = self.pw, self._lib
        T, ph, pwid, F, oc = self.T, self.ph, self.pw_, pw.features, pw.oc
        P = ph * pwid
        h1, w1, h2, w2 = 4 * ph, 4 * pwid, 2 * ph, 2 * pwid
        h4, w4 = (ph + 1) // 2, (pwid + 1) // 2
        R = self.rows
        esz = 2 if self.bf16 else 4
        # feature store: zero-initialised like the reference's (video_depth.py:228-231); slot R is the incoming frame
        shapes = [((h1 + 2) * (w1 + 2), oc[0]), ((h2 + 2) * (w2 + 2), oc[1]), (P, oc[2]), (h4 * w4, oc[3])]
        EB = self.enc_batch
        self.rings = [torch.zeros((R + EB, r, c), dtype=self.adt, device=self.device) for r, c in shapes]
        self.frame_bytes = [r * c * esz for r, c in shapes]
        width = max(T, R + EB, max(self.npreds))
        self.idx = torch.zeros((5, width), dtype=torch.int32, device=self.device)

        # ---- encode: enc_batch frames into the incoming slots R .. R + enc_batch - 1
        self.BT, self._prefix = EB, "enc."
        self._alias = {f"layer_{k + 1}": self.rings[k][R:R + EB].reshape(EB * shapes[k][0], shapes[k][1]) for k in range(4)}
        self.enc_taps = self._build_encoder()   # [EB * P, D] each: final norm applied, class token dropped
        n_encoder_ops = len(self.ops)
        self._reassemble(self.enc_taps)
        assert not self._alias
        self._end_stage("encode")
        # the reassemble half on its own, for callers that bring their own encoder features (`foward_single_image`)
        self.stages["reassemble"] = {"ops": self.stages["encode"]["ops"][n_encoder_ops:], "graph": None, "runs": 0}

        # ---- middle: the T-frame context (T-1 stored frames + the incoming one)
        self.BT, self._prefix = T, "mid."
        l3 = self.buf("l3in", T * P, oc[2])
        l4 = self.buf("l4in", T * h4 * w4, oc[3])
        ctx = self.idx[self.IDX_CTX].data_ptr()
        self.op(lib.vda_copy_frames, self.rings[2].data_ptr(), l3.data_ptr(), ctx, None, T, self.frame_bytes[2])
        self.op(lib.vda_copy_frames, self.rings[3].data_ptr(), l4.data_ptr(), ctx, None, T, self.frame_bytes[3])
        # motion module 3 is cut after its last temporal attention: what follows (output projection, feed-forward,
        # proj_out, residual) is per frame and only the predicted frames' path_3 is ever read (dpt_temporal.py:222-227),
        # so the tails finish it on their n frames instead of the middle stage on all T -- the same per-frame arithmetic
        mm3_parts = self._head_middle(l3, l4, skip_tmp_block, split_mm3=True)
        mm3_bytes = h2 * w2 * F * esz
        self._end_stage("middle")

        # ---- tails: the frames that get a depth map (dpt_temporal.py:187-192,222-227)
        self.depths = {}
        for n in self.npreds:
            self.BT, self._prefix = n, f"tail{n}."
            l1 = self.padded("l1in", h1, w1, oc[0])
            l2 = self.padded("l2in", h2, w2, oc[1])
            p3s = self.padded("p3in", h2, w2, F)
            pred = self.idx[self.IDX_PRED].data_ptr()
            self.op(lib.vda_copy_frames, self.rings[0].data_ptr(), l1.data_ptr(), pred, None, n, self.frame_bytes[0])
            self.op(lib.vda_copy_frames, self.rings[1].data_ptr(), l2.data_ptr(), pred, None, n, self.frame_bytes[1])
            parts = []
            for k, src_buf in enumerate(mm3_parts):
                dst_buf = self.buf(f"mm3in{k}", n * h2 * w2, F)
                self.op(lib.vda_copy_frames, src_buf.data_ptr(), dst_buf.data_ptr(), self.idx[self.IDX_P3].data_ptr(), None, n,
                        mm3_bytes)
                parts.append(dst_buf)
            self._mm_back(3, parts[0], parts[1], parts[2], (h2, w2), F, p3s, L.MAP_PADDED)
            self.depths[n] = self._head_tail(p3s, l1, l2)
            self._end_stage(f"tail{n}")

        # ---- commit: incoming slot idx[IDX_SRC][0] -> ring slot idx[IDX_COMMIT][0]
        src, dst = self.idx[self.IDX_SRC].data_ptr(), self.idx[self.IDX_COMMIT].data_ptr()
        for k in range(4):
            self.op(lib.vda_copy_frames, self.rings[k].data_ptr(), self.rings[k].data_ptr(), src, dst, 1, self.frame_bytes[k])
        self._end_stage("commit")
        self.BT, self._prefix = 1, ""
        self.depth = None

    # ------------------------------------------------------------------ feature store <-> the reference's NCHW tensors
    def ring_geometry(self):
        """(h, w, padded) of layer_1..4: layers 1 and 2 are stored with a zero border (they feed 3x3 convs)."""
        ph, pwid = self.ph, self.pw_
        return [(4 * ph, 4 * pwid, True), (2 * ph, 2 * pwid, True), (ph, pwid, False), ((ph + 1) // 2, (pwid + 1) // 2, False)]

    def store_features(self, k, slot0, t):
        """Write t [n, C, h, w] (the reference's layout of layer_{k+1}, dpt_temporal.py:101-130) into ring slots slot0.."""
        h, w, padded = self.ring_geometry()[k]
        n, C_ = t.shape[0], self.rings[k].shape[2]
        if tuple(t.shape[1:]) != (C_, h, w):
            raise ValueError(f"layer_{k + 1}: expected [n, {C_}, {h}, {w}], got {list(t.shape)}")
        src = t.to(device=self.device, dtype=self.adt).permute(0, 2, 3, 1)
        dst = self.rings[k][slot0:slot0 + n]
        if padded:
            dst.view(n, h + 2, w + 2, C_)[:, 1:-1, 1:-1].copy_(src)
        else:
            dst.view(n, h, w, C_).copy_(src)

    def load_features(self, k, slot):
        """Ring slot -> [1, C, h, w] in the build's activation type."""
        h, w, padded = self.ring_geometry()[k]
        C_ = self.rings[k].shape[2]
        v = self.rings[k][slot]
        v = v.view(h + 2, w + 2, C_)[1:-1, 1:-1] if padded else v.view(h, w, C_)
        return v.permute(2, 0, 1).unsqueeze(0).contiguous()

    # ------------------------------------------------------------------ execution
    def index_row(self, ctx=(), pred=(), p3=(), commit=(), src=()):
        """One step's index arrays as a host int32 [5, width] block (ring slots / context positions)."""
        row = torch.zeros(tuple(self.idx.shape), dtype=torch.int32)
        for r, vals in ((self.IDX_CTX, ctx), (self.IDX_PRED, pred), (self.IDX_P3, p3), (self.IDX_COMMIT, commit),
                        (self.IDX_SRC, src)):
            if len(vals):
                row[r, :len(vals)] = torch.as_tensor(list(vals), dtype=torch.int32)
        return row

    def load_indices(self, row_dev):
        """Make a step's index block current (device-to-device, stream ordered: no host buffer is reused in flight)."""
        self.idx.copy_(row_dev, non_blocking=True)

    def run_stage(self, name):
        """"encode" = encoder + reassemble; "reassemble" alone starts from `self.enc_taps` (written by the caller)."""
        st = self.stages[name]
        if st["graph"] is not None:
            st["graph"].replay()
            return
        self.ops = st["ops"]
        self._launch_ops()
        st["runs"] += 1
        if self.use_graph and st["runs"] == 2 and not torch.cuda.is_current_stream_capturing():
            g = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(g):
                    self._launch_ops()
                st["graph"] = g
            except Exception as e:  # capture is an optimisation
                self.use_graph = False
                _warn_graph_fallback(e)

    def launch(self):
        raise RuntimeError("StreamPlan runs stage by stage (run_stage)")

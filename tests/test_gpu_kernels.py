"""GPU parity tests of the individual kernels, called through the C ABI, against plain PyTorch references that
apply the same rounding points.  SIMT and memory-bound kernels (the tcgen05 kernels are in test_gpu_tc.py)."""
import ctypes as C

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from tests.util import bf16_close, stats

pytestmark = pytest.mark.gpu

dev = "cuda"


def _lib():
    from video_depth_anything_b200 import lib as L
    L.require_device("cuda")
    return L, L.load()


def R(t):
    return t.to(torch.bfloat16).float()


def run_gemm(L, lib, **kw):
    d = L.GemmDesc()
    keep = []
    for k, v in kw.items():
        if isinstance(v, torch.Tensor):
            keep.append(v)
            v = v.data_ptr()
        setattr(d, k, v)
    if not d.use_tc:
        ws = torch.empty(d.M * d.N * 4, dtype=torch.uint8, device=dev)
        d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
        keep.append(ws)
    L.gemm(d)
    torch.cuda.synchronize()


@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_gemm_simt_bias(mode):
    L, lib = _lib()
    torch.manual_seed(0)
    M, N, K = 300, 200, 136
    A, B, bias = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev) / K ** 0.5, torch.randn(N, device=dev)
    if mode == "bf16":
        A16, B16 = A.bfloat16(), B.bfloat16()
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        run_gemm(L, lib, A=A16, B=B16, lda=K, ldb=K, rowsA=M, a_dtype=L.BF16, b_dtype=L.BF16, M=M, N=N, K=K, taps=1, ksplit=1,
                 bias=bias, round_bf16=1, out_dtype=L.BF16, out=out, ldo=N)
        bf16_close("gemm_simt_bf16", out.float(), R(A16.float() @ B16.float().T + bias))
    else:
        out = torch.empty(M, N, device=dev)
        run_gemm(L, lib, A=A, B=B, lda=K, ldb=K, rowsA=M, a_dtype=L.F32, b_dtype=L.F32, M=M, N=N, K=K, taps=1, ksplit=1,
                 bias=bias, round_bf16=0, out_dtype=L.F32, out=out, ldo=N)
        ref = (A.double() @ B.double().T + bias.double()).float()
        s = stats("gemm_simt_f32", out, ref)
        assert s["max_rel_to_peak"] < 1e-5


def _pad_nhwc(x):  # [B,C,H,W] -> padded channels-last [B*(H+2)*(W+2), C]
    B, Cc, H, W = x.shape
    p = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1))
    return p.reshape(B * (H + 2) * (W + 2), Cc).contiguous()


def _unpad(t, B, H, W):
    return t.reshape(B, H + 2, W + 2, -1)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)


def test_conv3x3_simt_f32_all_maps():
    from video_depth_anything_b200.engine import _conv3x3_pack
    L, lib = _lib()
    torch.manual_seed(1)
    Bf, Ci, Co, H, W = 3, 24, 40, 9, 7
    x, w, b = torch.randn(Bf, Ci, H, W, device=dev), torch.randn(Co, Ci, 3, 3, device=dev) * 0.1, torch.randn(Co, device=dev)
    skip = torch.randn(Bf, Co, H, W, device=dev)
    xp, wp, sp = _pad_nhwc(x), _conv3x3_pack(w), _pad_nhwc(skip)
    Mp = xp.shape[0]
    common = dict(A=xp, B=wp, lda=Ci, ldb=9 * Ci, rowsA=Mp, a_dtype=L.F32, b_dtype=L.F32, M=Mp, N=Co, K=Ci, taps=9, ksplit=1,
                  gb=Bf, gh=H, gw=W, m_padded=1, bias=b, round_bf16=0, out_dtype=L.F32, ldo=Co)
    ref = F.conv2d(x, w, b, padding=1)
    # padded output with skip connection and relu copy
    out = torch.zeros(Mp, Co, device=dev)
    out_r = torch.zeros(Mp, Co, device=dev)
    run_gemm(L, lib, out=out, out_relu=out_r, res=sp, ldres=Co, **common)
    assert stats("conv_pad", _unpad(out, Bf, H, W), ref + skip)["max_rel_to_peak"] < 1e-5
    assert stats("conv_pad_relu", _unpad(out_r, Bf, H, W), F.relu(ref + skip))["max_rel_to_peak"] < 1e-5
    border = out.reshape(Bf, H + 2, W + 2, Co)
    assert float(border[:, 0].abs().max()) == 0 and float(border[:, :, 0].abs().max()) == 0  # borders stay zero
    # dense output
    outd = torch.zeros(Bf * H * W, Co, device=dev)
    run_gemm(L, lib, out=outd, out_map=L.MAP_DENSE, act=L.ACT_RELU, **common)
    assert stats("conv_dense_relu", outd.reshape(Bf, H, W, Co).permute(0, 3, 1, 2), F.relu(ref))["max_rel_to_peak"] < 1e-5
    # stride 2 (dpt.py:83-88) by sub-sampling the stride-1 result
    H2, W2 = (H + 1) // 2, (W + 1) // 2
    outs = torch.zeros(Bf * H2 * W2, Co, device=dev)
    run_gemm(L, lib, out=outs, out_map=L.MAP_STRIDE2, **common)
    ref2 = F.conv2d(x, w, b, stride=2, padding=1)
    assert ref2.shape[2:] == (H2, W2)
    assert stats("conv_s2", outs.reshape(Bf, H2, W2, Co).permute(0, 3, 1, 2), ref2)["max_rel_to_peak"] < 1e-5
    # fused head tail: relu -> 1x1 (Co=32 -> 1) -> relu
    w32, b32 = torch.randn(32, Ci, 3, 3, device=dev) * 0.1, torch.randn(32, device=dev)
    hw_, hb = torch.randn(32, device=dev), 0.3
    ho = torch.zeros(Bf * H * W, device=dev)
    c2 = dict(common)
    c2.update(B=_conv3x3_pack(w32), N=32, bias=b32)
    run_gemm(L, lib, out_map=L.MAP_DENSE, head_w=hw_, head_out=ho, head_b=hb, **c2)
    refh = F.relu(F.conv2d(F.relu(F.conv2d(x, w32, b32, padding=1)), hw_.reshape(1, 32, 1, 1), torch.tensor([hb], device=dev)))
    assert stats("conv_head", ho.reshape(Bf, 1, H, W), refh)["max_rel_to_peak"] < 1e-5


def test_convt_geglu_resid_simt_f32():
    from video_depth_anything_b200.engine import _convt_pack, _convt_stride, _geglu_interleave
    L, lib = _lib()
    torch.manual_seed(2)
    Bf, Ci, Co, H, W, k = 2, 48, 48, 5, 6, 4
    x, w, b = torch.randn(Bf, Ci, H, W, device=dev), torch.randn(Ci, Co, k, k, device=dev) * 0.1, torch.randn(Co, device=dev)
    xd = x.permute(0, 2, 3, 1).reshape(Bf * H * W, Ci).contiguous()
    out = torch.zeros(Bf * (H * k + 2) * (W * k + 2), Co, device=dev)
    run_gemm(L, lib, A=xd, B=_convt_pack(w), lda=Ci, ldb=Ci, rowsA=xd.shape[0], a_dtype=L.F32, b_dtype=L.F32, M=xd.shape[0],
             N=k * k * _convt_stride(Co), K=Ci, taps=1, ksplit=1, gb=Bf, gh=H, gw=W, bias=b, round_bf16=0, out_dtype=L.F32,
             out=out, ldo=Co, out_map=L.MAP_CONVT, ct_k=k, ct_cout=Co, ct_stride=_convt_stride(Co))
    # un-padded column blocks (ct_stride = 0 means ct_cout) stay supported
    out0 = torch.zeros_like(out)
    run_gemm(L, lib, A=xd, B=w.permute(2, 3, 1, 0).reshape(k * k * Co, Ci).contiguous(), lda=Ci, ldb=Ci, rowsA=xd.shape[0],
             a_dtype=L.F32, b_dtype=L.F32, M=xd.shape[0], N=k * k * Co, K=Ci, taps=1, ksplit=1, gb=Bf, gh=H, gw=W, bias=b,
             round_bf16=0, out_dtype=L.F32, out=out0, ldo=Co, out_map=L.MAP_CONVT, ct_k=k, ct_cout=Co)
    assert torch.equal(out0, out)
    ref = F.conv_transpose2d(x, w, b, stride=k)
    assert stats("convt", _unpad(out, Bf, H * k, W * k), ref)["max_rel_to_peak"] < 1e-5
    # GEGLU (motion_module/attention.py:382-384)
    M, Cc = 70, 64
    a, wg, bg = torch.randn(M, Cc, device=dev), torch.randn(8 * Cc, Cc, device=dev) * 0.1, torch.randn(8 * Cc, device=dev)
    og = torch.zeros(M, 4 * Cc, device=dev)
    run_gemm(L, lib, A=a, B=_geglu_interleave(wg), lda=Cc, ldb=Cc, rowsA=M, a_dtype=L.F32, b_dtype=L.F32, M=M, N=8 * Cc, K=Cc,
             taps=1, ksplit=1, bias=_geglu_interleave(bg), act=L.ACT_GEGLU, round_bf16=0, out_dtype=L.F32, out=og, ldo=4 * Cc)
    v, g = F.linear(a, wg, bg).chunk(2, -1)
    assert stats("geglu", og, v * F.gelu(g))["max_rel_to_peak"] < 1e-5
    # LayerScale + fp32 residual, and token-row assignment with pos-embed
    N = 96
    wr, br, gam = torch.randn(N, Cc, device=dev) * 0.1, torch.randn(N, device=dev), torch.rand(N, device=dev) + 0.5
    resid = torch.randn(M, N, device=dev)
    r0 = resid.clone()
    run_gemm(L, lib, A=a, B=wr, lda=Cc, ldb=Cc, rowsA=M, a_dtype=L.F32, b_dtype=L.F32, M=M, N=N, K=Cc, taps=1, ksplit=1, bias=br,
             round_bf16=0, out_dtype=L.F32, resid=resid, gamma=gam, ldresid=N)
    assert stats("resid_ls", resid, r0 + F.linear(a, wr, br) * gam)["max_rel_to_peak"] < 1e-5
    gh, gw = 5, 7  # 2 frames of 35 patches: rows m -> m + m/35 + 1
    pos = torch.randn(gh * gw + 1, N, device=dev)
    tok = torch.zeros(2 * (gh * gw + 1), N, device=dev)
    run_gemm(L, lib, A=a, B=wr, lda=Cc, ldb=Cc, rowsA=M, a_dtype=L.F32, b_dtype=L.F32, M=M, N=N, K=Cc, taps=1, ksplit=1, bias=br,
             gb=2, gh=gh, gw=gw, round_bf16=0, out_dtype=L.F32, resid=tok, pos=pos, ldresid=N, resid_map=L.MAP_TOKENS,
             resid_assign=1, pos_rows=gh * gw + 1)
    reft = (F.linear(a, wr, br).reshape(2, gh * gw, N) + pos[1:]).reshape(-1, N)
    assert stats("tokens", tok.reshape(2, gh * gw + 1, N)[:, 1:].reshape(-1, N), reft)["max_rel_to_peak"] < 1e-5


@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_layernorm_groupnorm(mode):
    L, lib = _lib()
    torch.manual_seed(3)
    st = L.stream_ptr()
    for D in (64, 192, 384, 1024):
        rows, grp = 3 * 11, 11
        x = torch.randn(rows, D, device=dev) * 2 + 0.5
        w, b = torch.rand(D, device=dev) + 0.5, torch.randn(D, device=dev)
        pe = torch.randn(4, D, device=dev)
        ydt = torch.bfloat16 if mode == "bf16" else torch.float32
        y = torch.zeros(rows - 3, D, device=dev, dtype=ydt)
        L.check(lib.vda_layernorm(x.data_ptr(), L.F32, y.data_ptr(), L.dt(y), w.data_ptr(), b.data_ptr(), 1e-6, rows, D, grp, 1,
                                  None, 0, 0, st), "ln")
        ref = F.layer_norm(x, (D,), w, b, 1e-6).reshape(3, grp, D)[:, 1:].reshape(-1, D)
        if mode == "bf16":
            bf16_close(f"ln_skip_{D}", y.float(), R(ref))
        else:
            assert stats(f"ln_skip_{D}", y, ref)["max_rel_to_peak"] < 2e-6
        # with positional encoding rows (motion_module.py:255-256): t = (row / pe_group) % pe_T
        xin = x.to(ydt)
        y2 = torch.zeros(rows, D, device=dev, dtype=ydt)
        L.check(lib.vda_layernorm(xin.data_ptr(), L.dt(xin), y2.data_ptr(), L.dt(y2), w.data_ptr(), b.data_ptr(), 1e-5, rows, D, 0,
                                  0, pe.data_ptr(), 3, 4, st), "ln_pe")
        t_idx = (torch.arange(rows, device=dev) // 3) % 4
        ref2 = F.layer_norm(xin.float(), (D,), w, b, 1e-5) + pe[t_idx]
        if mode == "bf16":
            bf16_close(f"ln_pe_{D}", y2.float(), R(ref2))
        else:
            assert stats(f"ln_pe_{D}", y2, ref2)["max_rel_to_peak"] < 2e-6
    for C_, hw in ((64, 50), (192, 37), (384, 9), (256, 130), (1024, 21)):
        frames = 3
        dt = torch.bfloat16 if mode == "bf16" else torch.float32
        x = (torch.randn(frames, hw, C_, device=dev) * 1.5 + 0.3).to(dt)
        w, b = torch.rand(C_, device=dev) + 0.5, torch.randn(C_, device=dev)
        y = torch.zeros_like(x)
        stt = torch.zeros(frames * 32 * 2 * 34, dtype=torch.float64, device=dev)
        L.check(lib.vda_groupnorm_tokens(x.data_ptr(), L.dt(x), y.data_ptr(), L.dt(y), w.data_ptr(), b.data_ptr(), 1e-6, frames, hw,
                                         C_, 32, stt.data_ptr(), st), "gn")
        ref = F.group_norm(x.float().permute(0, 2, 1), 32, w, b, 1e-6).permute(0, 2, 1)
        if mode == "bf16":
            bf16_close(f"gn_{C_}", y.float(), R(ref))
        else:
            assert stats(f"gn_{C_}", y, ref)["max_rel_to_peak"] < 5e-6


@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_bilinear_im2col_misc(mode):
    L, lib = _lib()
    torch.manual_seed(4)
    st = L.stream_ptr()
    dt = torch.bfloat16 if mode == "bf16" else torch.float32
    Bf, Cc = 2, 32
    # the last three sizes take the shared-memory row kernel in bf16 (ow >= 32): 2x, a non-integer ratio, a ragged last tile
    for (ih, iw, oh, ow) in ((5, 7, 10, 14), (9, 9, 19, 17), (12, 16, 21, 28), (19, 19, 37, 37), (37, 40, 74, 80), (20, 50, 33, 71)):
        x = torch.randn(Bf, Cc, ih, iw, device=dev).to(dt)
        xin = _pad_nhwc(x)
        out = torch.zeros(Bf * oh * ow, Cc, device=dev, dtype=dt)
        L.check(lib.vda_bilinear(xin.data_ptr(), L.dt(x), out.data_ptr(), L.dt(x), Bf, ih, iw, 1, oh, ow, 0, Cc, st), "bilinear")
        ref = F.interpolate(x, size=(oh, ow), mode="bilinear", align_corners=True)
        got = out.reshape(Bf, oh, ow, Cc).permute(0, 3, 1, 2)
        if mode == "bf16":
            bf16_close(f"bilinear_{oh}x{ow}", got.float(), ref.float())
        else:
            assert stats(f"bilinear_{oh}x{ow}", got, ref)["max_rel_to_peak"] < 2e-6
        outp = torch.zeros(Bf * (oh + 2) * (ow + 2), Cc, device=dev, dtype=dt)
        xd = x.permute(0, 2, 3, 1).reshape(-1, Cc).contiguous()
        L.check(lib.vda_bilinear(xd.data_ptr(), L.dt(x), outp.data_ptr(), L.dt(x), Bf, ih, iw, 0, oh, ow, 1, Cc, st), "bilinear")
        assert torch.equal(_unpad(outp, Bf, oh, ow), got)
        if mode == "bf16":  # bf16 in -> un-rounded fp32 out (what autocast's fp32 upsample produces)
            o32 = torch.zeros(Bf * oh * ow, Cc, device=dev)
            L.check(lib.vda_bilinear(xin.data_ptr(), L.BF16, o32.data_ptr(), L.F32, Bf, ih, iw, 1, oh, ow, 0, Cc, st), "bilinear")
            assert stats("bilinear_bf16_f32", o32.reshape(Bf, oh, ow, Cc).permute(0, 3, 1, 2), F.interpolate(
                x.float(), size=(oh, ow), mode="bilinear", align_corners=True))["max_rel_to_peak"] < 2e-6
    H, W = 28, 42
    x = torch.randn(3, 3, H, W, device=dev)
    A = torch.zeros(3 * 2 * 3, 640, device=dev, dtype=dt)
    L.check(lib.vda_im2col_patch(x.data_ptr(), A.data_ptr(), L.dt(A), 3, H, W, 640, st), "im2col")
    ref = F.unfold(x, 14, stride=14).transpose(1, 2).reshape(-1, 588)
    assert torch.equal(A[:, :588].float(), ref.to(dt).float()) and float(A[:, 588:].abs().max()) == 0
    # pos-embed bicubic (dinov2.py:190-198)
    D = 64
    pos = torch.randn(1370, D, device=dev)
    for (ph, pw) in ((3, 4), (37, 66), (20, 37)):
        out = torch.zeros(ph * pw + 1, D, device=dev)
        L.check(lib.vda_pos_embed_bicubic(pos.data_ptr(), out.data_ptr(), D, 37, ph, pw, (ph + 0.1) / 37.0, (pw + 0.1) / 37.0, st),
                "bicubic")
        ref = F.interpolate(pos[1:].reshape(1, 37, 37, D).permute(0, 3, 1, 2), scale_factor=((ph + 0.1) / 37.0, (pw + 0.1) / 37.0),
                            mode="bicubic", antialias=False)
        assert ref.shape[-2:] == (ph, pw)
        assert stats(f"bicubic_{ph}x{pw}", out[1:], ref.permute(0, 2, 3, 1).reshape(-1, D))["max_rel_to_peak"] < 1e-5
        assert torch.equal(out[0], pos[0])


@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_attention_simt(mode):
    L, lib = _lib()
    torch.manual_seed(5)
    st = L.stream_ptr()
    dt = torch.bfloat16 if mode == "bf16" else torch.float32
    frames, ntok, heads = 2, 150, 3
    D = heads * 64
    qkv = (torch.randn(frames * ntok, 3 * D, device=dev) * 1.2).to(dt)
    o = torch.zeros(frames * ntok, D, device=dev, dtype=dt)
    L.check(lib.vda_attention(qkv.data_ptr(), o.data_ptr(), L.dt(qkv), frames, ntok, heads, int(mode == "bf16"), 0, st), "attn")
    q, k, v = qkv.float().reshape(frames, ntok, 3, heads, 64).permute(2, 0, 3, 1, 4)
    if mode == "bf16":
        s = R((q * 0.125) @ k.transpose(-1, -2))
        ref = R(R(s.softmax(-1)) @ v)
        bf16_close("attn_simt_bf16", o.float().reshape(frames, ntok, heads, 64).permute(0, 2, 1, 3), ref, frac=5e-3)
    else:
        ref = ((q * 0.125) @ k.transpose(-1, -2)).softmax(-1) @ v
        assert stats("attn_simt_f32", o.reshape(frames, ntok, heads, 64).permute(0, 2, 1, 3), ref)["max_rel_to_peak"] < 1e-5
    # temporal attention (motion_module/attention.py:182-211)
    for (T, hw, C_) in ((4, 6, 64), (32, 5, 192), (32, 3, 1024), (16, 4, 256), (3, 7, 384), (32, 9, 256), (64, 3, 256), (64, 2, 1024),
                       (32, 7, 64), (16, 6, 64), (64, 3, 192), (32, 4, 384),
                       (1, 5, 64), (5, 4, 192), (10, 3, 256), (24, 5, 1024), (33, 2, 64), (50, 3, 256)):   # any T <= 64
        batch = 2 if T == 4 else 1
        qkv = (torch.randn(batch * T * hw, 3 * C_, device=dev)).to(dt)
        o = torch.zeros(batch * T * hw, C_, device=dev, dtype=dt)
        L.check(lib.vda_temporal_attention(qkv.data_ptr(), o.data_ptr(), L.dt(qkv), batch, T, hw, C_, 8, int(mode == "bf16"), st), "tattn")
        d = C_ // 8
        x = qkv.float().reshape(batch, T, hw, 3, 8, d).permute(3, 0, 2, 4, 1, 5)  # [3, b, p, h, T, d]
        q, k, v = x[0], x[1], x[2]
        s = (q @ k.transpose(-1, -2)) * d ** -0.5
        if mode == "bf16":
            ref = R(R(R(s).softmax(-1)) @ v)
        else:
            ref = s.softmax(-1) @ v
        got = o.float().reshape(batch, T, hw, 8, d).permute(0, 2, 3, 1, 4)
        if mode == "bf16":
            bf16_close(f"tattn_{T}_{C_}", got, ref, frac=5e-3)
        else:
            assert stats(f"tattn_{T}_{C_}", got, ref)["max_rel_to_peak"] < 1e-5


def test_driver_kernels():
    import cv2
    from oracle import driver as odrv
    from oracle.inputs import synthetic_video
    L, lib = _lib()
    st = L.stream_ptr()
    for (fh, fw, isz) in ((48, 64, 56), (72, 128, 70), (56, 56, 56), (37, 111, 56)):
        frames = synthetic_video(3, fh, fw, 7)
        e = odrv.effective_input_size(fh, fw, isz)
        nh, nw = odrv.network_size(fh, fw, e)
        ref = np.stack([odrv.preprocess_frame(f, e) for f in frames])
        fd = torch.from_numpy(frames).to(dev)
        out = torch.zeros(3, 3, nh, nw, device=dev)
        L.check(lib.vda_preprocess_u8(fd.data_ptr(), out.data_ptr(), 3, fh, fw, nh, nw, st), "pre")
        s = stats(f"preprocess_{fh}x{fw}", out.cpu(), torch.from_numpy(ref))
        assert s["max_abs"] < 2e-5
    d = torch.rand(4, 14, 28, device=dev) * 50
    for dt in (torch.float32, torch.bfloat16):
        din = d.to(dt)
        out = torch.zeros(4, 31, 45, device=dev)
        L.check(lib.vda_resize_depth(din.data_ptr(), L.dt(din), out.data_ptr(), 4, 14, 28, 31, 45, st), "resize_depth")
        ref = F.interpolate(din.float().unsqueeze(1), size=(31, 45), mode="bilinear", align_corners=True)[:, 0]
        assert stats("resize_depth", out, ref)["max_rel_to_peak"] < 1e-6
    p = torch.rand(2 * 33 * 47, device=dev) * 20
    t = p * 1.7 + 0.3 + torch.randn_like(p) * 0.01
    scratch = torch.zeros(5, dtype=torch.float64, device=dev)
    stt = torch.zeros(2, device=dev)
    L.check(lib.vda_scale_shift(p.data_ptr(), t.data_ptr(), p.numel(), scratch.data_ptr(), stt.data_ptr(), 3, st), "ls")
    s_ref, t_ref = odrv.compute_scale_and_shift(p.cpu().numpy(), t.cpu().numpy(), np.ones(p.numel()) == 1)
    got = stt.cpu().numpy()
    assert abs(got[0] - s_ref) < 1e-4 and abs(got[1] - t_ref) < 1e-3, (got, s_ref, t_ref)
    x = torch.randn(5 * 100, device=dev) * 10
    o = torch.zeros_like(x)
    L.check(lib.vda_scale_clamp(x.data_ptr(), o.data_ptr(), stt.data_ptr(), x.numel(), st), "sc")
    assert torch.equal(o, torch.clamp(x * stt[0] + stt[1], min=0))
    pre = torch.rand(8, 60, device=dev)
    post = torch.randn(8, 60, device=dev)
    refb = np.stack(odrv.blend_frames(list(pre.cpu().numpy()), [np.maximum(f * got[0] + got[1], 0).astype(np.float32)
                                                               for f in post.cpu().numpy()]))
    L.check(lib.vda_blend(pre.data_ptr(), post.data_ptr(), stt.data_ptr(), 8, 60, st), "blend")
    assert np.abs(pre.cpu().numpy() - refb).max() < 1e-6


def test_preprocess_720p_to_518x924():
    """BASELINE.json config 4's geometry: 720x1280 frames -> cv2.INTER_CUBIC to 518x924 -> normalise
    (util/transform.py:109-158) against the oracle's cv2 path, and on a ratio > 1.78 frame (374x1242 -> 280x924)."""
    from oracle import driver as odrv
    from oracle.inputs import synthetic_video
    L, lib = _lib()
    st = L.stream_ptr()
    for (fh, fw, want) in ((720, 1280, (518, 924)), (374, 1242, (280, 924))):
        frames = synthetic_video(2, fh, fw, 3)
        e = odrv.effective_input_size(fh, fw, 518)
        nh, nw = odrv.network_size(fh, fw, e)
        assert (nh, nw) == want
        ref = np.stack([odrv.preprocess_frame(f, e) for f in frames])
        fd = torch.from_numpy(frames).to(dev)
        out = torch.zeros(2, 3, nh, nw, device=dev)
        L.check(lib.vda_preprocess_u8(fd.data_ptr(), out.data_ptr(), 2, fh, fw, nh, nw, st), "pre")
        s = stats(f"preprocess_{fh}x{fw}", out.cpu(), torch.from_numpy(ref))
        # fp32 sums of four signed taps in another order than cv2's SIMD code: 4.0e-5 observed at the worst of 1.4 M values
        # in [-2.2, 2.7] (1.5e-5 of the range, mean 2e-6); the small cases of test_driver_kernels hold 2e-5
        assert s["max_abs"] < 1e-4 and s["mean_abs"] < 5e-6


def test_sharded_stitch_kernels():
    """vda_scale_shift_det (deterministic, grid-independent), vda_blend_raw (bit-identical to vda_scale_clamp + vda_blend),
    vda_copy_device, vda_im2col_patch_frames (bit-identical to vda_im2col_patch on gathered frames)."""
    from oracle import driver as odrv
    L, lib = _lib()
    st = L.stream_ptr()
    g = torch.Generator(device="cpu").manual_seed(5)
    for count in (2 * 33 * 47, 4096, 4097, 2 * 720 * 1280):
        p = (torch.rand(count, generator=g) * 20).to(dev)
        t = p * 1.7 + 0.3 + torch.randn(count, generator=g).to(dev) * 0.01
        nws = int(lib.vda_scale_shift_det_workspace_doubles(count))
        assert nws == ((count + 4095) // 4096) * 4
        outs = []
        for rep in range(3):   # poisoned scratch, repeated: bit-identical every time
            ws = torch.full((nws,), float("nan") if rep else 7.0, dtype=torch.float64, device=dev)
            stt = torch.zeros(2, device=dev)
            L.check(lib.vda_scale_shift_det(p.data_ptr(), t.data_ptr(), count, ws.data_ptr(), stt.data_ptr(), st), "ls_det")
            outs.append(stt.cpu().numpy().copy())
        assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
        s64, t64 = np.polyfit(p.double().cpu().numpy(), t.double().cpu().numpy(), 1)
        assert abs(outs[0][0] - s64) < 2e-6 and abs(outs[0][1] - t64) < 2e-5, (outs[0], s64, t64)
        s_ref, t_ref = odrv.compute_scale_and_shift(p.cpu().numpy(), t.cpu().numpy(), np.ones(count) == 1)
        assert abs(outs[0][0] - s_ref) < 1e-3 and abs(outs[0][1] - t_ref) < 1e-2   # the reference's fp32 numpy sums are the noisy side
    fe = 45 * 63   # not a multiple of 4: unaligned frames
    pre_raw = (torch.randn(8, fe, generator=g) * 3).to(dev)
    post_raw = (torch.randn(8, fe, generator=g) * 3).to(dev)
    st0 = torch.tensor([1.3, -0.2], device=dev)
    st1 = torch.tensor([0.7, 0.4], device=dev)
    for with_pre in (True, False):
        a = torch.empty_like(pre_raw)
        if with_pre:
            L.check(lib.vda_scale_clamp(pre_raw.data_ptr(), a.data_ptr(), st0.data_ptr(), pre_raw.numel(), st), "sc")
        else:
            L.check(lib.vda_copy_device(pre_raw.data_ptr(), a.data_ptr(), pre_raw.numel() * 4, st), "copy")
            assert torch.equal(a, pre_raw)
        L.check(lib.vda_blend(a.data_ptr(), post_raw.data_ptr(), st1.data_ptr(), 8, fe, st), "blend")
        b = torch.empty_like(pre_raw)
        L.check(lib.vda_blend_raw(pre_raw.data_ptr(), st0.data_ptr() if with_pre else None, post_raw.data_ptr(), st1.data_ptr(),
                                  b.data_ptr(), 8, fe, st), "blend_raw")
        assert torch.equal(a, b)
    for dt, H, W in ((torch.bfloat16, 28, 42), (torch.float32, 28, 42), (torch.bfloat16, 14, 14)):
        pool = torch.randn(7, 3, H, W, generator=g).to(dev)
        sel = [6, 0, 3, 3, 1]
        ptrs = torch.tensor([pool[i].data_ptr() for i in sel], dtype=torch.int64).to(dev)
        rows = len(sel) * (H // 14) * (W // 14)
        A0 = torch.zeros(rows, 640, dtype=dt, device=dev)
        A1 = torch.ones(rows, 640, dtype=dt, device=dev)
        xs = pool[sel].contiguous()
        L.check(lib.vda_im2col_patch(xs.data_ptr(), A0.data_ptr(), L.dt(A0), len(sel), H, W, 640, st), "im2col")
        L.check(lib.vda_im2col_patch_frames(ptrs.data_ptr(), A1.data_ptr(), L.dt(A1), len(sel), H, W, 640, st), "im2col_frames")
        assert torch.equal(A0, A1) and A0.float().abs().max().item() > 1


@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_rope_qk(mode):
    """vda_rope_qk against `apply_rotary_emb` restated with torch complex arithmetic (motion_module/attention.py:419-429)."""
    from oracle.model import rope_freqs_cis, rope_rotate
    from video_depth_anything_b200.engine import _rope_table
    L, lib = _lib()
    torch.manual_seed(3)
    b, T, hw, Cc = 2, 8, 37, 64
    rows = b * T * hw
    qkv = torch.randn(rows, 3 * Cc, device=dev)
    dt = torch.bfloat16 if mode == "bf16" else torch.float32
    buf = qkv.to(dt).contiguous()
    src = buf.clone()
    tab = _rope_table(Cc, 32, dev)
    L.check(lib.vda_rope_qk(buf.data_ptr(), L.dt(buf), tab.data_ptr(), rows, hw, T, Cc, int(mode == "bf16"), L.stream_ptr()),
            "vda_rope_qk")
    torch.cuda.synchronize()
    fc = rope_freqs_cis(Cc, 32)[:T].to(dev)
    # rows are ((b, t), p): bring q / k to the reference's [(b p), t, c] view, rotate, and back
    def ref_rot(x):
        y = x.reshape(b, T, hw, Cc).permute(0, 2, 1, 3).reshape(b * hw, T, Cc)
        y = rope_rotate(y, fc)
        return y.reshape(b, hw, T, Cc).permute(0, 2, 1, 3).reshape(rows, Cc)
    ref = torch.cat([ref_rot(src[:, :Cc]), ref_rot(src[:, Cc:2 * Cc]), src[:, 2 * Cc:]], 1)
    assert torch.equal(buf[:, 2 * Cc:], src[:, 2 * Cc:])  # v is untouched
    if mode == "bf16":
        bf16_close("rope_bf16", buf.float(), ref.float(), ulps=1.0, frac=1e-3)
    else:
        assert stats("rope_f32", buf, ref)["max_rel_to_peak"] < 1e-6


@pytest.mark.parametrize("case", ["s1_d80", "s2_d40"])
def test_eval_aligner_and_metrics(case):
    """the evaluation aligner + metrics on the device (vda_eval_fit / vda_eval_apply behind
    `evaluation.align_prediction`) vs the real reference's output (golden) and the oracle."""
    import os
    from oracle import driver as odrv
    from oracle.inputs import synthetic_eval_case
    from tests.util import GOLD
    from video_depth_anything_b200 import evaluation as pev
    g = np.load(os.path.join(GOLD, f"eval_{case}.npz"))
    seed, max_depth = [int(v) for v in g["meta"]]
    pred, gt, valid = synthetic_eval_case(seed)
    al, s, t, m = pev.align_prediction(pred, gt, valid, max_depth=float(max_depth), return_metrics=True)
    assert isinstance(al, np.ndarray) and al.dtype == np.float32 and al.shape == pred.shape
    # f64 normal equations vs LAPACK's f64 SVD lstsq
    assert np.allclose([s, t], g["scale_shift"], rtol=1e-8), (s, t, g["scale_shift"])
    ref = g["aligned"]
    e = np.abs(al[:, ::2, ::2] - ref) / np.maximum(np.abs(ref), 1e-3 * np.abs(ref).max())
    print(f"[eval {case}] scale {s:.6f} shift {t:.6f} aligned max rel-err {e.max():.3e}")
    assert e.max() <= 1e-6
    assert np.allclose([m[k] for k in sorted(m)], g["metrics"], rtol=1e-6, atol=1e-9), (m, g["metrics"])
    # device tensors in -> device tensor out, same numbers
    dp, dg, dv = (torch.from_numpy(x).to(dev) for x in (pred, gt, valid))
    al2, s2, t2 = pev.align_prediction(dp, dg, dv, max_depth=float(max_depth))
    # (the f64 block sums are combined with atomics: the last bits of the coefficients depend on the order)
    assert al2.is_cuda and np.allclose([s2, t2], [s, t], rtol=1e-12) and (al2.cpu().numpy() != al).mean() < 1e-4
    # with the oracle's coefficients the aligned map is the oracle's bit for bit (same float32 arithmetic)
    ev = pev.DeviceEvaluator(dev)
    _, so, to = odrv.align_prediction(pred, gt, valid, max_depth=float(max_depth))
    coef = torch.tensor([so, to], dtype=torch.float64, device=dev)
    al3, sums = ev.apply(dp, dg, dv.view(torch.uint8), coef, float(max_depth))
    oal = odrv.align_prediction(pred, gt, valid, max_depth=float(max_depth))[0].astype(np.float32)
    assert np.array_equal(al3.cpu().numpy(), oal)
    mo = odrv.eval_metrics(odrv.align_prediction(pred, gt, valid, max_depth=float(max_depth))[0], gt, valid)
    md = ev.metrics_from_sums(sums)
    assert all(abs(md[k] - mo[k]) <= 1e-9 * max(1.0, abs(mo[k])) for k in mo), (md, mo)

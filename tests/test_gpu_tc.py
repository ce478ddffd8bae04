"""GPU parity tests of the tcgen05 / TMEM / TMA kernels (GEMM + conv-as-GEMM with fused epilogues, flash
attention) through the C ABI, against fp32 PyTorch references with the same bf16 rounding points."""
import pytest
import torch
import torch.nn.functional as F

from tests.test_gpu_kernels import R, _lib, _pad_nhwc, _unpad, dev, run_gemm
from tests.util import bf16_close, stats

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (300, 256, 128), (1000, 64, 192), (777, 48, 384), (4500, 3072, 1024),
                                   (20000, 1024, 4096), (260, 32, 64), (129, 384, 1152)])
def test_gemm_tc_bias(M, N, K):
    L, lib = _lib()
    torch.manual_seed(M + N + K)
    A = torch.randn(M, K, device=dev).bfloat16()
    B = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device=dev)
    out = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
    run_gemm(L, lib, A=A, B=B, lda=K, ldb=K, rowsA=M, a_dtype=L.BF16, b_dtype=L.BF16, M=M, N=N, K=K, taps=1, ksplit=1, bias=bias,
             round_bf16=1, out_dtype=L.BF16, out=out, ldo=N, use_tc=1)
    ref = R(A.float() @ B.float().T + bias)
    bf16_close(f"gemm_tc_{M}x{N}x{K}", out.float(), ref)


def test_gemm_tc_epilogues():
    from video_depth_anything_b200.engine import _convt_pack, _convt_stride, _geglu_interleave
    L, lib = _lib()
    torch.manual_seed(11)
    M, K, N = 1500, 256, 512
    A = torch.randn(M, K, device=dev).bfloat16()
    Wt = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device=dev)
    base = dict(A=A, lda=K, ldb=K, rowsA=M, a_dtype=L.BF16, b_dtype=L.BF16, M=M, K=K, taps=1, ksplit=1, round_bf16=1,
                out_dtype=L.BF16, use_tc=1)
    lin = R(A.float() @ Wt.float().T + bias)
    # GELU (mlp.py:37-38 under autocast: bf16 linear output, bf16 gelu output)
    out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
    run_gemm(L, lib, B=Wt, N=N, bias=bias, act=L.ACT_GELU, out=out, ldo=N, **base)
    bf16_close("tc_gelu", out.float(), R(F.gelu(lin)))
    # LayerScale + fp32 residual (block.py:105-106)
    gam = torch.rand(N, device=dev) + 0.5
    resid = torch.randn(M, N, device=dev)
    r0 = resid.clone()
    run_gemm(L, lib, B=Wt, N=N, bias=bias, resid=resid, gamma=gam, ldresid=N, **base)
    s = stats("tc_resid", resid, r0 + lin * gam)
    assert s["max_rel_to_peak"] < 2e-2 and s["mean_abs"] < 2e-3
    # bf16 residual in place + relu copy
    res = torch.randn(M, N, device=dev).bfloat16()
    o2 = res.clone()
    orl = torch.zeros_like(o2)
    run_gemm(L, lib, B=Wt, N=N, bias=bias, out=o2, out_relu=orl, ldo=N, res=o2, ldres=N, **base)
    bf16_close("tc_res_inplace", o2.float(), R(lin + res.float()))
    assert torch.equal(orl, torch.relu(o2))
    # GEGLU
    Wg = (torch.randn(2 * N, K, device=dev) / K ** 0.5).bfloat16()
    bg = torch.randn(2 * N, device=dev)
    og = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
    run_gemm(L, lib, B=_geglu_interleave(Wg), N=2 * N, bias=_geglu_interleave(bg), act=L.ACT_GEGLU, out=og, ldo=N, **base)
    v, g = R(A.float() @ Wg.float().T + bg).chunk(2, -1)
    bf16_close("tc_geglu", og.float(), R(v * R(F.gelu(g))))
    # conv-transpose k = s scatter into a padded map; channel counts that are not multiples of 64 (vits: 48, 96) travel in
    # column blocks padded to 64 / 128 (ct_stride) so the coalesced epilogue applies
    for (Bf, H, W_, Ci, Co, k) in ((3, 9, 11, 96, 96, 2), (2, 7, 5, 48, 48, 4), (2, 6, 7, 256, 256, 4)):
        x = torch.randn(Bf, Ci, H, W_, device=dev).bfloat16()
        w = (torch.randn(Ci, Co, k, k, device=dev) * 0.1).bfloat16()
        b = torch.randn(Co, device=dev)
        xd = x.permute(0, 2, 3, 1).reshape(-1, Ci).contiguous()
        out = torch.zeros(Bf * (H * k + 2) * (W_ * k + 2), Co, device=dev, dtype=torch.bfloat16)
        run_gemm(L, lib, A=xd, B=_convt_pack(w), lda=Ci, ldb=Ci, rowsA=xd.shape[0], a_dtype=L.BF16, b_dtype=L.BF16, M=xd.shape[0],
                 N=k * k * _convt_stride(Co), K=Ci, taps=1, ksplit=1, gb=Bf, gh=H, gw=W_, bias=b, round_bf16=1, out_dtype=L.BF16,
                 out=out, ldo=Co, out_map=L.MAP_CONVT, ct_k=k, ct_cout=Co, ct_stride=_convt_stride(Co), use_tc=1)
        ref = R(F.conv_transpose2d(x.float(), w.float(), b, stride=k))
        bf16_close(f"tc_convt_{Co}_{k}", _unpad(out, Bf, H * k, W_ * k).float(), ref)
        g = out.reshape(Bf, H * k + 2, W_ * k + 2, Co)
        assert float(g[:, 0].abs().max()) == 0 and float(g[:, :, -1].abs().max()) == 0  # borders untouched


@pytest.mark.parametrize("Ci,Co,H,W", [(64, 64, 12, 10), (48, 64, 20, 17), (256, 128, 30, 30), (1024, 256, 19, 19),
                                        (256, 128, 61, 57), (128, 96, 70, 45), (64, 64, 40, 44), (64, 32, 50, 37),
                                        (192, 64, 37, 37)])
def test_conv3x3_tc(Ci, Co, H, W):
    from video_depth_anything_b200.engine import _conv3x3_pack
    L, lib = _lib()
    torch.manual_seed(Ci + Co)
    Bf = 3
    x = torch.randn(Bf, Ci, H, W, device=dev).bfloat16()
    w = (torch.randn(Co, Ci, 3, 3, device=dev) / (9 * Ci) ** 0.5).bfloat16()
    b = torch.randn(Co, device=dev)
    skip = torch.randn(Bf, Co, H, W, device=dev).bfloat16()
    xp, sp = _pad_nhwc(x), _pad_nhwc(skip)
    Mp = xp.shape[0]
    out = torch.zeros(Mp, Co, device=dev, dtype=torch.bfloat16)
    common = dict(A=xp, B=_conv3x3_pack(w), lda=Ci, ldb=9 * Ci, rowsA=Mp, a_dtype=L.BF16, b_dtype=L.BF16, M=Mp, N=Co, K=Ci, taps=9,
                  ksplit=1, gb=Bf, gh=H, gw=W, m_padded=1, bias=b, round_bf16=1, out_dtype=L.BF16, ldo=Co, use_tc=1)
    run_gemm(L, lib, out=out, res=sp, ldres=Co, **common)
    conv = R(F.conv2d(x.float(), w.float(), b, padding=1))
    bf16_close(f"tc_conv_{Ci}_{Co}", _unpad(out, Bf, H, W).float(), R(conv + skip.float()))
    g = out.reshape(Bf, H + 2, W + 2, Co)
    assert float(g[:, 0].abs().max()) == 0 and float(g[:, :, -1].abs().max()) == 0
    H2, W2 = (H + 1) // 2, (W + 1) // 2
    outs = torch.zeros(Bf * H2 * W2, Co, device=dev, dtype=torch.bfloat16)
    run_gemm(L, lib, out=outs, out_map=L.MAP_STRIDE2, **common)
    ref2 = R(F.conv2d(x.float(), w.float(), b, stride=2, padding=1))
    bf16_close(f"tc_conv_s2_{Ci}_{Co}", outs.reshape(Bf, H2, W2, Co).permute(0, 3, 1, 2).float(), ref2)


def test_output_head_tc_tf32():
    """output_conv2 (dpt.py:118-124) in the bf16 build: fp32 activations and weights through tcgen05 kind::tf32 (what
    cuDNN does for an fp32 conv by default), fused ReLU -> 1x1 -> ReLU, fp32 result."""
    from video_depth_anything_b200.engine import _conv3x3_pack
    L, lib = _lib()
    torch.manual_seed(12)
    for Ci in (32, 128):
        Bf, H, W = 2, 28, 42
        x = torch.randn(Bf, Ci, H, W, device=dev)
        w = torch.randn(32, Ci, 3, 3, device=dev) / (9 * Ci) ** 0.5
        b, hw_, hb = torch.randn(32, device=dev) * 0.1, torch.randn(32, device=dev).abs(), 0.05
        xp = _pad_nhwc(x)
        Mp = xp.shape[0]
        ho = torch.zeros(Bf * H * W, device=dev)
        run_gemm(L, lib, A=xp, B=_conv3x3_pack(w), lda=Ci, ldb=9 * Ci, rowsA=Mp, a_dtype=L.F32, b_dtype=L.F32, M=Mp, N=32, K=Ci,
                 taps=9, ksplit=1, gb=Bf, gh=H, gw=W, m_padded=1, bias=b, round_bf16=1, out_dtype=L.BF16, out_map=L.MAP_DENSE,
                 head_w=hw_, head_out=ho, head_b=hb, use_tc=1)
        ref = F.relu(F.conv2d(F.relu(F.conv2d(x.double(), w.double(), b.double(), padding=1)), hw_.double().reshape(1, 32, 1, 1),
                              torch.tensor([hb], device=dev, dtype=torch.float64)))
        s = stats(f"tc_head_tf32_{Ci}", ho.reshape(Bf, 1, H, W), ref)
        assert s["max_rel_to_peak"] < 2e-3 and s["mean_abs"] < 3e-4 * s["ref_peak"]  # tf32: 10-bit mantissa operands


@pytest.mark.parametrize("frames,ntok,heads", [(1, 128, 1), (2, 150, 3), (2, 1370, 6), (1, 2443, 2)])
def test_flash_attention_tc(frames, ntok, heads):
    L, lib = _lib()
    torch.manual_seed(ntok)
    st = L.stream_ptr()
    D = heads * 64
    qkv = (torch.randn(frames * ntok, 3 * D, device=dev) * 1.3).bfloat16()
    o = torch.full((frames * ntok, D), float("nan"), device=dev, dtype=torch.bfloat16)
    L.check(lib.vda_attention(qkv.data_ptr(), o.data_ptr(), L.BF16, frames, ntok, heads, 1, 1, st), "attn_tc")
    torch.cuda.synchronize()
    q, k, v = qkv.float().reshape(frames, ntok, 3, heads, 64).permute(2, 0, 3, 1, 4)
    s = R((q * 0.125) @ k.transpose(-1, -2))
    ref = R(R(s.softmax(-1)) @ v)
    got = o.float().reshape(frames, ntok, heads, 64).permute(0, 2, 1, 3)
    # The flash kernel rounds un-normalised probabilities exp(s - running max) to bf16 where the reference rounds the
    # normalised softmax: two independent bf16 roundings of P.  So the gate is statistical: its error against exact
    # fp32 attention must be of the same size as the reference's own bf16 error against exact attention.
    exact = s.softmax(-1) @ v
    e_ref = stats(f"ref_bf16_vs_exact_{ntok}", ref, exact)
    e_got = stats(f"flash_tc_vs_exact_{ntok}", got, exact)
    assert e_got["mean_abs"] <= 1.5 * e_ref["mean_abs"] and e_got["max_abs"] <= 3.0 * e_ref["max_abs"]
    d = stats(f"flash_tc_vs_ref_{ntok}", got, ref)
    assert d["mean_abs"] <= 2.0 * e_ref["mean_abs"] and d["max_abs"] <= 4.0 * e_ref["max_abs"]
    # and it must agree equally well with the SIMT kernel that follows the reference's rounding exactly
    o2 = torch.zeros_like(o)
    L.check(lib.vda_attention(qkv.data_ptr(), o2.data_ptr(), L.BF16, frames, ntok, heads, 1, 0, st), "attn_simt")
    d2 = stats(f"flash_tc_vs_simt_{ntok}", o.float(), o2.float())
    assert d2["mean_abs"] <= 2.0 * e_ref["mean_abs"]


@pytest.mark.parametrize("Ci,ih,iw,oh,ow,frames", [(32, 16, 24, 28, 42, 2), (128, 32, 48, 56, 84, 3), (64, 9, 7, 21, 13, 1),
                                                   (128, 40, 40, 70, 70, 2)])
def test_output_head_fused(Ci, ih, iw, oh, ow, frames):
    """vda_output_head: bilinear (align_corners) -> 3x3 conv (TF32) -> ReLU -> 1x1 -> ReLU without materialising the
    upsampled fp32 map; must agree with fp64 torch to TF32 accuracy and with the unfused bilinear + vda_gemm pair."""
    from video_depth_anything_b200.engine import _conv3x3_pack
    L, lib = _lib()
    torch.manual_seed(Ci + oh)
    x = torch.randn(frames, Ci, ih, iw, device=dev).bfloat16()
    w = torch.randn(32, Ci, 3, 3, device=dev) / (9 * Ci) ** 0.5
    b, hw_, hb = torch.randn(32, device=dev) * 0.1, torch.randn(32, device=dev).abs(), 0.05
    xd = x.permute(0, 2, 3, 1).reshape(-1, Ci).contiguous()
    wp = _conv3x3_pack(w)
    depth = torch.full((frames, oh, ow), float("nan"), device=dev)
    L.check(lib.vda_output_head(xd.data_ptr(), wp.data_ptr(), b.data_ptr(), hw_.data_ptr(), hb, depth.data_ptr(), frames, ih, iw,
                                Ci, oh, ow, L.stream_ptr()), "vda_output_head")
    torch.cuda.synchronize()
    up = F.interpolate(x.double(), (oh, ow), mode="bilinear", align_corners=True)
    ref = F.relu(F.conv2d(F.relu(F.conv2d(up, w.double(), b.double(), padding=1)), hw_.double().reshape(1, 32, 1, 1),
                          torch.tensor([hb], device=dev, dtype=torch.float64)))
    s = stats(f"head_fused_{Ci}_{oh}x{ow}", depth.reshape(frames, 1, oh, ow), ref)
    assert s["max_rel_to_peak"] < 2e-3 and s["mean_abs"] < 3e-4 * s["ref_peak"]  # tf32: 10-bit mantissa operands
    # unfused pair through the C ABI
    upb = torch.zeros(frames * (oh + 2) * (ow + 2), Ci, device=dev)
    L.check(lib.vda_bilinear(xd.data_ptr(), L.BF16, upb.data_ptr(), L.F32, frames, ih, iw, 0, oh, ow, 1, Ci, L.stream_ptr()), "bl")
    ho = torch.zeros(frames * oh * ow, device=dev)
    run_gemm(L, lib, A=upb, B=wp, lda=Ci, ldb=9 * Ci, rowsA=upb.shape[0], a_dtype=L.F32, b_dtype=L.F32, M=upb.shape[0], N=32, K=Ci,
             taps=9, ksplit=1, gb=frames, gh=oh, gw=ow, m_padded=1, bias=b, round_bf16=1, out_dtype=L.BF16, out_map=L.MAP_DENSE,
             head_w=hw_, head_out=ho, head_b=hb, use_tc=1)
    s2 = stats("head_fused_vs_unfused", depth.reshape(-1), ho)
    assert s2["max_rel_to_peak"] < 1e-3


@pytest.mark.parametrize("Ci,ih,iw,oh,ow,frames", [(64, 9, 7, 21, 13, 1), (128, 32, 48, 56, 84, 3), (128, 40, 40, 70, 70, 2),
                                                   (192, 12, 20, 30, 37, 2)])
def test_output_head_fused_f16(Ci, ih, iw, oh, ow, frames):
    """vda_output_head_f16 (fp16 operands, 64 channels per chunk): same bar against fp64 torch as the TF32 form, and no
    farther from it than the TF32 form is; saturates instead of overflowing."""
    from video_depth_anything_b200.engine import _conv3x3_pack
    L, lib = _lib()
    torch.manual_seed(Ci + oh)
    x = torch.randn(frames, Ci, ih, iw, device=dev).bfloat16()
    w = torch.randn(32, Ci, 3, 3, device=dev) / (9 * Ci) ** 0.5
    b, hw_, hb = torch.randn(32, device=dev) * 0.1, torch.randn(32, device=dev).abs(), 0.05
    xd = x.permute(0, 2, 3, 1).reshape(-1, Ci).contiguous()
    wp = _conv3x3_pack(w)
    wh = wp.to(torch.float16).contiguous()
    depth = torch.full((frames, oh, ow), float("nan"), device=dev)
    L.check(lib.vda_output_head_f16(xd.data_ptr(), wh.data_ptr(), b.data_ptr(), hw_.data_ptr(), hb, depth.data_ptr(), frames,
                                    ih, iw, Ci, oh, ow, L.stream_ptr()), "vda_output_head_f16")
    d32 = torch.empty_like(depth)
    L.check(lib.vda_output_head(xd.data_ptr(), wp.data_ptr(), b.data_ptr(), hw_.data_ptr(), hb, d32.data_ptr(), frames, ih, iw,
                                Ci, oh, ow, L.stream_ptr()), "vda_output_head")
    torch.cuda.synchronize()
    up = F.interpolate(x.double(), (oh, ow), mode="bilinear", align_corners=True)
    ref = F.relu(F.conv2d(F.relu(F.conv2d(up, w.double(), b.double(), padding=1)), hw_.double().reshape(1, 32, 1, 1),
                          torch.tensor([hb], device=dev, dtype=torch.float64)))
    s = stats(f"head_f16_{Ci}_{oh}x{ow}", depth.reshape(frames, 1, oh, ow), ref)
    s32 = stats(f"head_tf32_{Ci}_{oh}x{ow}", d32.reshape(frames, 1, oh, ow), ref)
    assert s["max_rel_to_peak"] < 2e-3 and s["mean_abs"] < 3e-4 * s["ref_peak"]
    assert s["mean_abs"] <= 1.1 * s32["mean_abs"]  # round-to-nearest operands: at least as close as the truncating TF32 path
    # out-of-range activations saturate (no inf / nan in the depth map)
    xd2 = (xd.float() * 3e4).bfloat16()
    L.check(lib.vda_output_head_f16(xd2.data_ptr(), wh.data_ptr(), b.data_ptr(), hw_.data_ptr(), hb, depth.data_ptr(), frames,
                                    ih, iw, Ci, oh, ow, L.stream_ptr()), "vda_output_head_f16")
    torch.cuda.synchronize()
    assert torch.isfinite(depth).all()


@pytest.mark.parametrize("Ci,ih,iw,oh,ow,frames", [(128, 32, 48, 56, 84, 3), (64, 9, 7, 21, 13, 1)])
def test_output_head_guarded_falls_back_to_tf32_outside_fp16_range(Ci, ih, iw, oh, ow, frames):
    """vda_output_head_guarded: in range it IS the fp16-operand kernel (bit-identical, flag clear); with activations beyond
    fp16's range (where vda_output_head_f16 alone saturates at +-65504 and returns wrong values) the device-side guard
    selects the TF32 kernel -- bit-identical to vda_output_head, and right against fp64 torch.  No host synchronisation
    decides anything."""
    from video_depth_anything_b200.engine import _conv3x3_pack
    L, lib = _lib()
    torch.manual_seed(Ci + oh + 1)
    x = torch.randn(frames, Ci, ih, iw, device=dev).bfloat16()
    w = torch.randn(32, Ci, 3, 3, device=dev) / (9 * Ci) ** 0.5
    b, hw_, hb = torch.randn(32, device=dev) * 0.1, torch.randn(32, device=dev).abs(), 0.05
    wp = _conv3x3_pack(w)
    wh = wp.to(torch.float16).contiguous()
    flag = torch.full((1,), 7, dtype=torch.int32, device=dev)
    st = L.stream_ptr()

    def run(xin):
        xd = xin.permute(0, 2, 3, 1).reshape(-1, Ci).contiguous()
        outs = [torch.full((frames, oh, ow), float("nan"), device=dev) for _ in range(3)]
        L.check(lib.vda_output_head_guarded(xd.data_ptr(), wh.data_ptr(), wp.data_ptr(), b.data_ptr(), hw_.data_ptr(), hb,
                                            outs[0].data_ptr(), frames, ih, iw, Ci, oh, ow, flag.data_ptr(), st), "guarded")
        L.check(lib.vda_output_head_f16(xd.data_ptr(), wh.data_ptr(), b.data_ptr(), hw_.data_ptr(), hb, outs[1].data_ptr(), frames,
                                        ih, iw, Ci, oh, ow, st), "f16")
        L.check(lib.vda_output_head(xd.data_ptr(), wp.data_ptr(), b.data_ptr(), hw_.data_ptr(), hb, outs[2].data_ptr(), frames, ih,
                                    iw, Ci, oh, ow, st), "tf32")
        torch.cuda.synchronize()
        up = F.interpolate(xin.double(), (oh, ow), mode="bilinear", align_corners=True)
        ref = F.relu(F.conv2d(F.relu(F.conv2d(up, w.double(), b.double(), padding=1)), hw_.double().reshape(1, 32, 1, 1),
                              torch.tensor([hb], device=dev, dtype=torch.float64)))
        return outs, ref.reshape(frames, oh, ow), int(flag.item())

    (g, h, t), ref, fl = run(x)
    assert fl == 0 and torch.equal(g, h) and not torch.equal(g, t)
    # one element far outside fp16's range, the rest as before: the whole launch is recomputed in TF32
    x2 = x.clone()
    x2[0, 3, ih // 2, iw // 2] = 3.0e5
    (g, h, t), ref, fl = run(x2)
    assert fl == 1 and torch.equal(g, t)
    s = stats("head_guarded_out_of_range", g, ref)
    assert s["max_rel_to_peak"] < 2e-3
    assert stats("head_f16_saturated", h, ref)["max_rel_to_peak"] > 0.05   # what the unguarded fp16 form would have returned
    # the same with every activation large, with an infinity, and with a NaN: always the TF32 kernel's result
    (g, h, t), ref, fl = run((x.float() * 3e4).bfloat16())
    assert fl == 1 and torch.equal(g, t) and stats("head_guarded_all_large", g, ref)["max_rel_to_peak"] < 2e-3
    x3 = x.clone()
    x3[0, 0, 0, 0] = float("inf")
    (g, h, t), _, fl = run(x3)
    assert fl == 1 and torch.equal(g.nan_to_num(1.0, 2.0, 3.0), t.nan_to_num(1.0, 2.0, 3.0))
    # back in range: the flag is cleared by the call itself
    (g, h, t), ref, fl = run(x)
    assert fl == 0 and torch.equal(g, h)


@pytest.mark.parametrize("kind", ["store", "gelu", "resid"])
def test_pair_gemm_is_bit_reproducible_back_to_back(kind):
    """The pipeline hand-offs of the pair GEMM (relaxed cluster arrive on the accumulator-free barrier, elected-lane issue,
    double-buffered TMEM accumulators) carry no data: a hand-off that fired early would let an MMA overwrite an accumulator
    that is still being read, which shows as outputs that change from launch to launch.  40 back-to-back launches on an
    encoder-sized problem (27 tiles per CTA pair) must reproduce the first launch bit for bit."""
    L, lib = _lib()
    torch.manual_seed(5)
    M, N, K = 43840, 1024, 1024
    A = torch.randn(M, K, device=dev).bfloat16()
    B = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device=dev)
    base = dict(A=A, B=B, lda=K, ldb=K, rowsA=M, a_dtype=L.BF16, b_dtype=L.BF16, M=M, N=N, K=K, taps=1, ksplit=1, bias=bias,
                round_bf16=1, out_dtype=L.BF16, use_tc=1)
    gam = torch.rand(N, device=dev) + 0.5
    first = None
    for it in range(40):
        if kind == "resid":
            out = torch.zeros(M, N, device=dev)   # resid += gamma * R(acc + bias): one reduction per element and launch
            run_gemm(L, lib, resid=out, gamma=gam, ldresid=N, **base)
        else:
            out = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
            run_gemm(L, lib, out=out, ldo=N, act=L.ACT_GELU if kind == "gelu" else L.ACT_NONE, **base)
        if first is None:
            first = out
            ref = A[:512].float() @ B.float().T + bias
            if kind == "gelu":
                ref = F.gelu(R(ref))
            got = out[:512].float() / gam if kind == "resid" else out[:512].float()
            assert (got - ref).abs().max().item() < 0.05 * ref.abs().max().item()
        else:
            assert torch.equal(out, first), (kind, it)


def test_flash_attention_is_bit_reproducible_back_to_back():
    L, lib = _lib()
    torch.manual_seed(6)
    frames, ntok, heads = 32, 1370, 16
    qkv = (torch.randn(frames * ntok, 3 * heads * 64, device=dev) * 1.3).bfloat16()
    first = None
    for it in range(20):
        o = torch.full((frames * ntok, heads * 64), float("nan"), device=dev, dtype=torch.bfloat16)
        L.check(lib.vda_attention(qkv.data_ptr(), o.data_ptr(), L.BF16, frames, ntok, heads, 1, 1, L.stream_ptr()), "attn_tc")
        if first is None:
            first = o
            assert torch.isfinite(o.float()).all()
        else:
            assert torch.equal(o, first), it

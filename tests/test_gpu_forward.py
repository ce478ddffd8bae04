"""End-to-end parity of `VideoDepthAnything.forward` and `infer_video_depth` on the GPU:
  * fp32 build  vs the committed goldens (real reference, CPU fp32) and vs the oracle run on the GPU in strict fp32:
                per-pixel rel-err <= 1e-3 (BASELINE.json north_star)
  * bf16 build  vs the oracle under `torch.autocast('cuda', bfloat16)` (same autocast decisions as the reference):
                per-pixel rel-err <= 1e-2 and |AbsRel| <= 1e-3
Gate metric: |d - ref| / max(|ref|, 1e-3 max|ref|)."""
import os

import numpy as np
import pytest
import torch

from oracle import driver as odrv
from oracle import model as omodel
from oracle.inputs import synthetic_clip, synthetic_video
from oracle.weights import MODEL_CONFIGS, make_state_dict
from tests.util import GOLD, abs_rel, bf16_gate, rel_err_map, stats

pytestmark = pytest.mark.gpu
dev = "cuda"

_models = {}


def get_model(enc, pe="ape"):
    from video_depth_anything_b200.video_depth import VideoDepthAnything
    if (enc, pe) not in _models:
        m = VideoDepthAnything(**MODEL_CONFIGS[enc], pe=pe)
        sd = make_state_dict(enc, pe=pe)
        m.load_state_dict(sd, strict=True)
        _models[(enc, pe)] = (m.to(dev).eval(), {k: v.to(dev) for k, v in sd.items()})
    return _models[(enc, pe)]


def oracle_fp32(sd, x, enc, skip=False):
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            return omodel.forward(sd, x, enc, skip)
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old


def oracle_bf16(sd, x, enc, skip=False):
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        return omodel.forward(sd, x, enc, skip)


@pytest.mark.parametrize("case", ["vits_T4_42x56", "vits_T32_28x28_skip", "vitl_T3_28x42", "vits_T2_518x518",
                                  "vits_T8_42x56_rope"])
def test_fp32_build_matches_reference_golden(case):
    g = np.load(os.path.join(GOLD, f"forward_{case}.npz"))
    T, H, W, seed, sub, skip = [int(v) for v in g["meta"]]
    enc = "vitl" if case.startswith("vitl") else "vits"
    m, sd = get_model(enc, "rope" if case.endswith("rope") else "ape")
    x = synthetic_clip(1, T, H, W, seed).to(dev)
    with torch.no_grad():
        d = m(x, bool(skip))
    assert d.shape == (1, T, H, W) and d.dtype == torch.float32
    ref = torch.from_numpy(g["depth"]).to(dev)
    e = rel_err_map(d[0, :, ::sub, ::sub], ref)
    stats(f"fp32_vs_golden_{case}", d[0, :, ::sub, ::sub], ref)
    print(f"[{case}] max rel-err {e.max().item():.3e}")
    assert e.max().item() <= 1e-3
    o = oracle_fp32(sd, x, enc, bool(skip))
    assert rel_err_map(d, o).max().item() <= 1e-3


@pytest.mark.parametrize("enc,T,H,W,mode", [("vits", 4, 42, 56, "simt"), ("vits", 4, 42, 56, "tc"), ("vits", 32, 70, 98, "tc"),
                                             ("vits", 32, 70, 98, "tc+simt_attn"), ("vitl", 3, 28, 42, "tc"),
                                             ("vitl", 8, 56, 84, "tc"), ("vitl", 8, 56, 84, "tc+simt_attn"),
                                             ("vits", 8, 154, 210, "tc"), ("vits", 8, 154, 210, "simt")])
def test_bf16_build_matches_autocast_oracle(enc, T, H, W, mode):
    m, sd = get_model(enc)
    x = synthetic_clip(1, T, H, W, 21).to(dev)
    m.force_simt = mode == "simt"
    m.attn_simt = mode.endswith("simt_attn")
    try:
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            d = m(x)
    finally:
        m.force_simt = False
        m.attn_simt = False
    # under CUDA autocast the head's bilinear upsample is an fp32 op, so the reference returns fp32 (torch 2.11)
    assert d.dtype == torch.float32 and d.shape == (1, T, H, W)
    ref = oracle_bf16(sd, x, enc)
    assert ref.dtype == torch.float32
    f32 = oracle_fp32(sd, x, enc)
    bf16_gate(f"bf16 {enc} T{T} {H}x{W} {mode}", d.float(), ref.float(), f32)


@pytest.mark.parametrize("enc,T,H,W", [("vits", 32, 70, 98), ("vitl", 8, 56, 84)])
def test_bf16_build_rope_matches_autocast_oracle(enc, T, H, W):
    """pe='rope' (SURVEY section 8f rank 3): q/k rotated by vda_rope_qk between the fused qkv GEMM and the attention."""
    m, sd = get_model(enc, "rope")
    x = synthetic_clip(1, T, H, W, 23).to(dev)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        d = m(x)
    ref, f32 = oracle_bf16(sd, x, enc), oracle_fp32(sd, x, enc)
    bf16_gate(f"bf16 rope {enc} T{T} {H}x{W}", d.float(), ref.float(), f32)


def test_intermediate_taps_fp32():
    """Tap-level parity: encoder taps and motion-module outputs, fp32 build vs oracle."""
    enc, T, H, W = "vits", 4, 42, 56
    m, sd = get_model(enc)
    x = synthetic_clip(1, T, H, W, 5).to(dev)
    plan = m.plan(1, T, H, W, torch.device("cuda", torch.cuda.current_device()), False, keep=("tokens0", "mm0", "mm1"))
    plan.run(x.flatten(0, 1))
    torch.cuda.synchronize()
    col = {}
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            omodel.forward(sd, x, enc, collect=col)
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
    ph, pw = H // 14, W // 14
    for i in range(4):
        s = stats(f"tap{i}", plan.bufs[f"tap{i}"].reshape(T, ph * pw, -1), col[f"tap{i}"])
        assert s["max_rel_to_peak"] < 1e-4

    def unpad(t, h, w):
        return t.reshape(T, h + 2, w + 2, -1)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)

    assert stats("mm0", unpad(plan.bufs["mm0"], ph, pw), col["mm0"])["max_rel_to_peak"] < 1e-4
    assert stats("mm1", unpad(plan.bufs["mm1"], (ph + 1) // 2, (pw + 1) // 2), col["mm1"])["max_rel_to_peak"] < 1e-4
    assert stats("layer_1", unpad(plan.bufs["layer_1"], 4 * ph, 4 * pw), col["layer_1"])["max_rel_to_peak"] < 1e-4
    assert stats("path_1", unpad(plan.bufs["path_1"], 8 * ph, 8 * pw), col["path_1"])["max_rel_to_peak"] < 1e-4


def test_motion_module_forward_hooks_fire():
    """The reference's diagnostic harness hooks `model.head.motion_modules[i]` (Test_singleImage_processing.py:132-142):
    hooks registered on the drop-in fire once per forward with (input, output) of the temporal module as [B, C, T, h, w]."""
    enc, T, H, W = "vits", 4, 42, 56
    m, sd = get_model(enc)
    x = synthetic_clip(1, T, H, W, 6).to(dev)
    seen = {}
    handles = [mm.register_forward_hook(lambda mod, inp, out, i=i: seen.__setitem__(i, (inp, out)))
               for i, mm in enumerate(m.head.motion_modules)]
    try:
        with torch.no_grad():
            m(x)
    finally:
        for h in handles:
            h.remove()
    col = {}
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            omodel.forward(sd, x, enc, collect=col)
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
    assert sorted(seen) == [0, 1, 2, 3]
    for i, (kin, kout) in enumerate([("layer_3", "mm0"), ("layer_4", "mm1"), ("path_4_pre", "path_4"), ("path_3_pre", "path_3")]):
        (xin, a, b), out = seen[i]
        assert a is None and b is None and xin.shape == out.shape and xin.dim() == 5 and xin.shape[2] == T
        ref_in = col[kin].unflatten(0, (1, T)).permute(0, 2, 1, 3, 4)
        ref_out = col[kout].unflatten(0, (1, T)).permute(0, 2, 1, 3, 4)
        assert stats(f"hook{i}_in", xin, ref_in)["max_rel_to_peak"] < 1e-4
        assert stats(f"hook{i}_out", out, ref_out)["max_rel_to_peak"] < 1e-4
    with torch.no_grad():
        m(x)   # hooks removed: nothing fires, nothing breaks
    assert len(seen) == 4


def test_forward_contract():
    """Drop-in behaviours: batch > 1, shape assertion, determinism, permutation sensitivity."""
    m, sd = get_model("vits")
    x = synthetic_clip(2, 4, 28, 42, 9).to(dev)
    with torch.no_grad():
        d = m(x)
        d2 = m(x)
        assert torch.equal(d, d2)
        o = oracle_fp32(sd, x, "vits")
        assert rel_err_map(d, o).max().item() <= 1e-3
        # temporal modules are live (proj_out is not zero in the recipe): frame order matters
        perm = [1, 0, 2, 3]
        dp = m(x[:, perm])
        assert not torch.allclose(dp[:, perm], d, rtol=1e-3, atol=1e-3)
        with pytest.raises(AssertionError):
            m(torch.zeros(1, 2, 3, 30, 42, device=dev))
    from video_depth_anything_b200 import lib
    with pytest.raises(lib.VdaError):
        m(torch.zeros(1, 2, 3, 28, 28))  # CPU tensors are refused: there is no fallback path


def test_graph_replay_matches_launch_loop():
    """The plan captures its launch list into a CUDA graph after two eager runs: replays must reproduce the eager
    result bit for bit (fp32 and bf16 builds), on fresh inputs too."""
    m, sd = get_model("vits")
    xa, xb = synthetic_clip(1, 4, 42, 56, 21).to(dev), synthetic_clip(1, 4, 42, 56, 22).to(dev)
    for bf16 in (False, True):
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            plan = m.plan(1, 4, 42, 56, xa.device, bf16)  # the plan forward() uses (keyed by the tensor's device)
            plan._graph, plan._eager_runs, plan.use_graph = None, 0, True
            ea, eb = m(xa), m(xb)          # two eager runs; the second one also captures
            assert plan._graph is not None
            ga, gb = m(xa), m(xb)          # replays
            assert torch.equal(ea, ga) and torch.equal(eb, gb)
            assert not torch.equal(ga, gb)


@pytest.mark.parametrize("case", ["vits_n50_48x64", "vits_n23_40x90"])
def test_driver_fp32_matches_reference_golden(case):
    g = np.load(os.path.join(GOLD, f"driver_{case}.npz"))
    n, fh, fw, isz, seed = [int(v) for v in g["meta"]]
    m, sd = get_model("vits")
    frames = synthetic_video(n, fh, fw, seed)
    out, fps = m.infer_video_depth(frames, 24, input_size=isz, device="cuda", fp32=True)
    assert fps == 24 and out.shape == (n, fh, fw) and out.dtype == np.float32
    e = rel_err_map(torch.from_numpy(out), torch.from_numpy(g["depth"]))
    print(f"[driver {case}] max rel-err {e.max().item():.3e}")
    assert e.max().item() <= 1e-3


def test_driver_bf16_matches_autocast_oracle():
    n, fh, fw, isz = 50, 48, 64, 56
    m, sd = get_model("vits")
    frames = synthetic_video(n, fh, fw, 5)
    out, _ = m.infer_video_depth(frames, 24, input_size=isz, device="cuda", fp32=False)
    ref, _ = odrv.infer_video_depth(frames, 24, lambda x: omodel.forward(sd, x, "vits"), input_size=isz, device="cuda",
                                    fp32=False, autocast_dtype=torch.bfloat16)
    ref32, _ = odrv.infer_video_depth(frames, 24, lambda x: oracle_fp32(sd, x, "vits"), input_size=isz, device="cuda", fp32=True)
    bf16_gate("driver bf16", torch.from_numpy(out), torch.from_numpy(ref), torch.from_numpy(ref32))


STREAM_CASES = ["vits_n40_T8_k25_align", "vits_n40_T8_k03_noalign", "vits_n50_T32_k212_align"]


def _stream_meta(g):
    n, fh, fw, isz, seed, T, align = [int(v) for v in g["meta"][:7]]
    return n, fh, fw, isz, seed, T, bool(align), [int(v) for v in g["meta"][7:]]


def _oracle_stream(sd, frames, isz, T, kf, align, fp32, fwd32=False):
    def wrap(fn):
        if not fwd32:
            return fn
        def g(*a):
            old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
            torch.backends.cuda.matmul.allow_tf32 = False
            torch.backends.cudnn.allow_tf32 = False
            try:
                return fn(*a)
            finally:
                torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
        return g

    def reas(x):
        return omodel.reassemble(sd, omodel.encoder_taps(sd, x.flatten(0, 1), "vits"), x.shape[-2] // 14, x.shape[-1] // 14)

    def step(x, mf, pi):
        return omodel.forward_single_image(sd, x, "vits", mf, pi, T)

    return odrv.infere_single_image(frames, 24, wrap(step), wrap(reas), input_size=isz, device="cuda", fp32=fp32,
                                    inference_length=T, keyframe_list=kf, align_each_new_frame=align,
                                    autocast_dtype=torch.bfloat16)[0]


@pytest.mark.parametrize("case", STREAM_CASES)
def test_stream_fp32_matches_reference_golden(case):
    """Streaming mode: `infere_single_image` fp32 build vs the real reference's output."""
    g = np.load(os.path.join(GOLD, f"stream_{case}.npz"))
    n, fh, fw, isz, seed, T, align, kf = _stream_meta(g)
    m, sd = get_model("vits")
    frames = synthetic_video(n, fh, fw, seed)
    out, fps = m.infere_single_image(frames, 24, input_size=isz, device="cuda", fp32=True, inference_length=T,
                                     keyframe_list=kf, align_each_new_frame=align)
    assert fps == 24 and out.shape == g["depth"].shape and out.dtype == np.float32
    e = rel_err_map(torch.from_numpy(out), torch.from_numpy(g["depth"]))
    print(f"[stream {case}] max rel-err {e.max().item():.3e}")
    assert e.max().item() <= 1e-3
    out2, _ = m.infere_single_image(frames, 24, input_size=isz, device="cuda", fp32=True, inference_length=T,
                                    keyframe_list=kf, align_each_new_frame=align)
    assert np.array_equal(out, out2)  # second call: CUDA-graph replays of the stages, store re-zeroed


@pytest.mark.parametrize("T,kf,align", [(8, [2, 5], True), (32, [0, 12], False)])
def test_stream_bf16_matches_autocast_oracle(T, kf, align):
    n, fh, fw, isz = 44, 56, 84, 56
    m, sd = get_model("vits")
    frames = synthetic_video(n, fh, fw, 11)
    out, _ = m.infere_single_image(frames, 24, input_size=isz, device="cuda", fp32=False, inference_length=T,
                                   keyframe_list=kf, align_each_new_frame=align)
    ref = _oracle_stream(sd, frames, isz, T, kf, align, fp32=False)
    ref32 = _oracle_stream(sd, frames, isz, T, kf, align, fp32=True, fwd32=True)
    assert out.shape == ref.shape
    bf16_gate(f"stream bf16 T{T} kf{kf} align{int(align)}", torch.from_numpy(out), torch.from_numpy(ref), torch.from_numpy(ref32))


def test_stream_contract():
    """Failure modes of the reference's streaming entry point (video_depth.py:321-327, dpt_temporal.py:188-189)."""
    m, sd = get_model("vits")
    frames = synthetic_video(34, 28, 42, 9)
    with pytest.raises(IndexError):   # the reference's own defaults (keyframe_list=[0, 12], align) index past the context
        m.infere_single_image(frames, 24, input_size=28, device="cuda", fp32=True)
    with pytest.raises(NotImplementedError):
        m.infere_single_image(frames, 24, input_size=28, device="cuda", fp32=True, warmup=False)
    with pytest.raises(ValueError):   # fewer frames than the context: nothing to stack
        m.infere_single_image(frames[:20], 24, input_size=28, device="cuda", fp32=True, align_each_new_frame=False)


@pytest.mark.parametrize("enc,T,H,W,pred,skip", [("vits", 8, 42, 56, None, False), ("vits", 6, 56, 42, [0, 3], False),
                                                 ("vitl", 4, 28, 42, [1], True)])
def test_tensor_level_single_image_entry_points(enc, T, H, W, pred, skip):
    """`forward_single_image` (video_depth.py:67-91), `head.foward_single_image` (dpt_temporal.py:133-260) and
    `head.get_motion_features` (:101-130) with caller-held context tensors, fp32 build against the oracle in strict fp32;
    then the bf16 build under autocast against the oracle's autocast run through the usual gate."""
    m, sd = get_model(enc)
    ph, pw = H // 14, W // 14
    frames = synthetic_clip(1, T, H, W, 23).to(dev)[0]   # T-1 context frames + the new one
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            taps_ctx = omodel.encoder_taps(sd, frames[:T - 1], enc)
            feats_ref = omodel.reassemble(sd, taps_ctx, ph, pw)
            x_new = frames[T - 1:].unsqueeze(0)
            d_ref, new_ref = omodel.forward_single_image(sd, x_new, enc, tuple(feats_ref), None if pred is None else list(pred), T, skip)
            taps_new = omodel.encoder_taps(sd, frames[T - 1:], enc)
            # ---- get_motion_features: the context store from encoder features, (tokens, class token) pairs as the reference passes
            feats = m.head.get_motion_features([(t, t[:, 0]) for t in taps_ctx], ph, pw)
            for k in range(4):
                assert feats[k].shape == feats_ref[k].shape
                assert ((feats[k] - feats_ref[k]).abs().max() / feats_ref[k].abs().max()).item() <= 1e-4, k   # signed maps: to peak
            # ---- model level
            pred_arg = None if pred is None else list(pred)
            d, new = m.forward_single_image(x_new, feats, pred_arg, inference_length=T, skip_tmp_block=skip)
            assert pred_arg == (None if pred is None else list(pred))   # the reference appends and pops; the list comes back intact
            n = 1 if pred is None else len(pred) + 1
            assert d.shape == d_ref.shape == (1, n, H, W) and d.dtype == torch.float32
            assert rel_err_map(d, d_ref).max().item() <= 1e-3
            for k in range(4):
                assert new[k].shape == new_ref[k].shape
                assert ((new[k] - new_ref[k]).abs().max() / new_ref[k].abs().max()).item() <= 1e-4, k
            # ---- head level, from the oracle's encoder features of the new frame
            out = m.head.foward_single_image([(t, t[:, 0]) for t in taps_new], ph, pw, T, feats_ref, pred_arg, skip_tmp_block=skip)
            assert len(out) == 5 and out[0].shape == (n, 1, H, W)
            assert rel_err_map(out[0].squeeze(1), d_ref[0]).max().item() <= 1e-3
            # ---- a context of the wrong length fails like the reference's unflatten
            with pytest.raises(RuntimeError):
                m.forward_single_image(x_new, tuple(f[:-1] for f in feats), None, inference_length=T)
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        feats16 = omodel.reassemble(sd, omodel.encoder_taps(sd, frames[:T - 1], enc), ph, pw)
        d16_ref, _ = omodel.forward_single_image(sd, x_new, enc, tuple(feats16), None if pred is None else list(pred), T, skip)
        d16, new16 = m.forward_single_image(x_new, tuple(feats16), None if pred is None else list(pred), inference_length=T,
                                            skip_tmp_block=skip)
    assert new16[0].dtype == torch.bfloat16 and d16.dtype == torch.float32
    bf16_gate(f"single_image_{enc}_T{T}", d16, d16_ref, d_ref)


@pytest.mark.parametrize("enc,T,ph,pw", [("vits", 16, 5, 7), ("vits", 64, 4, 6), ("vitl", 16, 3, 4), ("vitl", 64, 2, 3)])
def test_isolated_head_frames_16_64(enc, T, ph, pw):
    """BASELINE.json config 5: `DPTHeadTemporal.forward` alone on random taps, 16 and 64 frames (64 needs
    num_frames=64 at construction), fp32 build vs the oracle and bf16 build vs the autocast oracle."""
    from video_depth_anything_b200.video_depth import VideoDepthAnything
    nf = max(T, 32)
    m = VideoDepthAnything(**MODEL_CONFIGS[enc], num_frames=nf)
    sd = make_state_dict(enc, num_frames=nf)
    m.load_state_dict(sd, strict=True)
    m = m.to(dev).eval()
    sd = {k: v.to(dev) for k, v in sd.items()}
    D = m.pretrained.embed_dim
    g = torch.Generator().manual_seed(31)
    taps = [torch.randn(T, ph * pw, D, generator=g).to(dev) for _ in range(4)]
    d = torch.device("cuda", torch.cuda.current_device())
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            ref32 = omodel.head_forward(sd, taps, ph, pw, T)[:, 0]
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        ref16 = omodel.head_forward(sd, taps, ph, pw, T)[:, 0]
    outs = {}
    for bf16 in (False, True):
        plan = m.head_plan(1, T, 14 * ph, 14 * pw, d, bf16)
        for t_dst, t_src in zip(plan.taps, taps):
            t_dst.copy_(t_src.reshape(t_dst.shape))
        plan.launch()
        torch.cuda.synchronize()
        outs[bf16] = plan.depth.clone()
    e = rel_err_map(outs[False], ref32).max().item()
    print(f"[head {enc} T{T}] fp32 max rel-err {e:.3e}")
    assert e <= 1e-3
    bf16_gate(f"head bf16 {enc} T{T}", outs[True].float(), ref16.float(), ref32)

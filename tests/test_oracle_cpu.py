"""CPU suite: the oracle against the committed goldens (generated from the real reference by
scripts/make_goldens.py), the host-side driver logic, and the C-ABI surface of the built library."""
import ctypes
import json
import os
import re

import numpy as np
import pytest
import torch

from oracle import driver as odrv
from oracle import model as omodel
from oracle.inputs import synthetic_clip, synthetic_video
from oracle.weights import make_state_dict, param_spec
from tests.util import GOLD



def _load(name):
    return np.load(os.path.join(GOLD, name))


@pytest.mark.parametrize("enc", ["vits", "vitl"])
def test_param_spec_matches_reference_keys(enc):
    with open(os.path.join(GOLD, f"state_dict_keys_{enc}.json")) as f:
        ref = json.load(f)
    spec = param_spec(enc)
    assert list(spec.keys()) == list(ref.keys())
    assert all(list(spec[k]) == ref[k] for k in ref)
    assert len(spec) == {"vits": 351, "vitl": 519}[enc]


@pytest.mark.parametrize("enc", ["vits", "vitl"])
def test_product_module_has_reference_state_dict(enc):
    from oracle.weights import MODEL_CONFIGS
    from video_depth_anything_b200.video_depth import VideoDepthAnything

    with open(os.path.join(GOLD, f"state_dict_keys_{enc}.json")) as f:
        ref = json.load(f)
    m = VideoDepthAnything(**MODEL_CONFIGS[enc])
    sd = m.state_dict()
    assert list(sd.keys()) == list(ref.keys())
    assert all(list(sd[k].shape) == ref[k] for k in ref)
    m.load_state_dict(make_state_dict(enc), strict=True)  # the drop-in contract (run.py:80)
    # zero-initialised proj_out and pe buffers like the reference
    m2 = VideoDepthAnything(**MODEL_CONFIGS[enc])
    assert float(m2.head.motion_modules[0].temporal_transformer.proj_out.weight.abs().max()) == 0.0
    from oracle.weights import sinusoidal_pe
    pe = m2.state_dict()["head.motion_modules.2.temporal_transformer.transformer_blocks.0.attention_blocks.1.pos_encoder.pe"]
    assert torch.equal(pe, sinusoidal_pe(32, pe.shape[-1]))


def test_rope_model_has_no_pe_buffers():
    """pe='rope' (video_depth.py:44 -> motion_module.py:238-242): freqs_cis is an attribute, not a buffer, so the state
    dict is the 'ape' one minus the 8 `pos_encoder.pe` entries; the drop-in loads it with strict=True."""
    from oracle.weights import MODEL_CONFIGS
    from video_depth_anything_b200.video_depth import VideoDepthAnything

    spec = param_spec("vits", pe="rope")
    assert len(spec) == 351 - 8 and not any(k.endswith(".pe") for k in spec)
    m = VideoDepthAnything(**MODEL_CONFIGS["vits"], pe="rope")
    assert list(m.state_dict().keys()) == list(spec.keys())
    m.load_state_dict(make_state_dict("vits", pe="rope"), strict=True)
    with pytest.raises(NotImplementedError):
        VideoDepthAnything(**MODEL_CONFIGS["vits"], pe="alibi")


@pytest.mark.parametrize("case", ["vits_T4_42x56", "vits_T32_28x28_skip", "vitl_T3_28x42", "vits_T2_518x518",
                                  "vits_T8_42x56_rope"])
def test_oracle_forward_matches_reference_golden(case):
    g = _load(f"forward_{case}.npz")
    T, H, W, seed, sub, skip = [int(v) for v in g["meta"]]
    enc = "vitl" if case.startswith("vitl") else "vits"
    sd = make_state_dict(enc, pe="rope" if case.endswith("rope") else "ape")
    x = synthetic_clip(1, T, H, W, seed)
    col = {}
    with torch.no_grad():
        d = omodel.forward(sd, x, enc, bool(skip), collect=col)[0, :, ::sub, ::sub].numpy()
    ref = g["depth"]
    assert d.shape == ref.shape
    assert np.abs(d - ref).max() <= 2e-5 * np.abs(ref).max()
    for i in range(4):
        t = col[f"tap{i}"]
        tt = t[:, :: max(1, t.shape[1] // 64), :: max(1, t.shape[2] // 48)].numpy()
        assert np.abs(tt - g[f"tap{i}"]).max() <= 2e-5 * max(1.0, np.abs(g[f"tap{i}"]).max())


@pytest.mark.parametrize("case", ["vits_n50_48x64", "vits_n23_40x90"])
def test_oracle_driver_matches_reference_golden(case):
    g = _load(f"driver_{case}.npz")
    n, fh, fw, isz, seed = [int(v) for v in g["meta"]]
    sd = make_state_dict("vits")
    frames = synthetic_video(n, fh, fw, seed)
    out, fps, raw, st = odrv.infer_video_depth(frames, 24, lambda x: omodel.forward(sd, x, "vits"), input_size=isz,
                                              device="cpu", fp32=True, return_raw=True)
    assert fps == 24 and out.shape == (n, fh, fw) and out.dtype == np.float32
    assert np.abs(out - g["depth"]).max() <= 2e-5 * np.abs(g["depth"]).max()
    assert np.allclose(np.asarray(st), g["scale_shift"], rtol=1e-4, atol=1e-4)


STREAM_CASES = ["vits_n40_T8_k25_align", "vits_n40_T8_k03_noalign", "vits_n50_T32_k212_align"]


def stream_meta(g):
    n, fh, fw, isz, seed, T, align = [int(v) for v in g["meta"][:7]]
    return n, fh, fw, isz, seed, T, bool(align), [int(v) for v in g["meta"][7:]]


@pytest.mark.parametrize("case", STREAM_CASES)
def test_oracle_stream_matches_reference_golden(case):
    """`infere_single_image` (video_depth.py:93-327): oracle restatement vs the real reference's output."""
    g = _load(f"stream_{case}.npz")
    n, fh, fw, isz, seed, T, align, kf = stream_meta(g)
    sd = make_state_dict("vits")
    frames = synthetic_video(n, fh, fw, seed)

    def reas(x):
        return omodel.reassemble(sd, omodel.encoder_taps(sd, x.flatten(0, 1), "vits"), x.shape[-2] // 14, x.shape[-1] // 14)

    out, fps = odrv.infere_single_image(frames, 24, lambda x, mf, pi: omodel.forward_single_image(sd, x, "vits", mf, pi, T),
                                        reas, input_size=isz, inference_length=T, keyframe_list=kf, align_each_new_frame=align)
    assert fps == 24 and out.shape == g["depth"].shape == ((n - 1) if align else (n - T + 1), fh, fw)
    assert np.abs(out - g["depth"]).max() <= 2e-5 * np.abs(g["depth"]).max()


def test_stream_schedule_and_feature_store():
    """Host logic of the streaming mode: the product's schedule equals the oracle's (pinned to the reference through the
    goldens above), and the slot table reproduces the reference's shifting store (video_depth.py:281-296)."""
    from video_depth_anything_b200 import stream as pstream

    for T, kf in [(32, [0, 12]), (32, [2, 12]), (8, [2, 5]), (8, [0, 3]), (16, [3, 7, 20]), (8, [9, 3]), (32, [5, 40])]:
        assert pstream.stream_schedule(T, kf) == odrv.stream_schedule(T, kf)
    with pytest.raises(AssertionError):
        pstream.stream_schedule(8, [3, 3])
    ctx, pos = pstream.stream_schedule(32, [0, 12])
    assert ctx[0][:3] == [0, 32, 2] and pos[0] == [0, 1, 20]  # the default key row 32 does not exist yet: IndexError upstream
    rows = 11
    store = pstream.FeatureStore(rows)
    literal = [-1] * rows          # the reference's tensor, holding frame numbers
    phys = [-1] * rows             # what each physical slot holds
    move = [i for i in range(rows) if i != 1]
    for i in range(40):
        phys[store.write(i)] = i
        if i < rows:
            literal[i] = i
        else:
            literal[:-1] = [literal[j] for j in move]
            literal[-1] = i
        assert [phys[s] for s in store.slot] == literal


@pytest.mark.parametrize("case", ["s1_d80", "s2_d40"])
def test_oracle_eval_aligner_matches_reference_golden(case):
    """utils/align.py:202-236 + utils/metrics.py CSV row: oracle restatement vs the real reference's output."""
    from oracle.inputs import synthetic_eval_case
    g = _load(f"eval_{case}.npz")
    seed, max_depth = [int(v) for v in g["meta"]]
    pred, gt, valid = synthetic_eval_case(seed)
    al, s, t = odrv.align_prediction(pred, gt, valid, max_depth=float(max_depth))
    assert np.array_equal(al[:, ::2, ::2].astype(np.float32), g["aligned"])
    assert np.allclose([s, t], g["scale_shift"], rtol=1e-6)
    m = odrv.eval_metrics(al, gt, valid)
    assert np.allclose([m[k] for k in sorted(m)], g["metrics"], rtol=1e-6)


def test_window_index_map_golden_and_product():
    from video_depth_anything_b200 import driver as pdrv

    with open(os.path.join(GOLD, "window_index_map.json")) as f:
        maps = json.load(f)
    for n, ref in maps.items():
        assert odrv.window_index_map(int(n)).tolist() == ref
        assert pdrv.window_index_map(int(n)).tolist() == ref
    # SURVEY section 3.1 worked example
    assert odrv.window_index_map(70)[2].tolist() == [0, 34] + list(range(46, 54)) + list(range(54, 70)) + [69] * 6
    assert odrv.window_index_map(1024).shape == (47, 32) and odrv.window_index_map(2048).shape == (94, 32)


def test_product_host_geometry_matches_oracle():
    from video_depth_anything_b200 import driver as pdrv

    for fh, fw, isz in [(518, 518, 518), (720, 1280, 518), (374, 1242, 518), (48, 64, 56), (40, 90, 56), (37, 111, 56)]:
        e = odrv.effective_input_size(fh, fw, isz)
        assert pdrv.effective_input_size(fh, fw, isz) == e
        assert pdrv.network_size(fh, fw, e) == odrv.network_size(fh, fw, e)
    assert odrv.network_size(720, 1280, 518) == (518, 924)
    assert odrv.network_size(374, 1242, odrv.effective_input_size(374, 1242, 518)) == (280, 924)
    for nw, world in [(47, 8), (94, 8), (3, 8), (1, 2), (5, 2)]:
        owned = [pdrv.rank_windows(nw, world, r) for r in range(world)]
        assert sorted(k for o in owned for k in o) == list(range(nw))
        assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1


def test_scale_shift_known_answer():
    rng = np.random.default_rng(0)
    p = rng.random((2 * 40, 60)).astype(np.float32) * 10
    s, t = odrv.compute_scale_and_shift(p, p * 1.7 + 0.3, np.ones_like(p) == 1)
    assert abs(s - 1.7) < 1e-3 and abs(t - 0.3) < 1e-2
    s, t = odrv.compute_scale_and_shift(np.ones((4, 4), np.float32), np.ones((4, 4), np.float32), np.ones((4, 4)) == 1)
    assert (s, t) == (1, 0)  # singular system keeps the identity (utils/util.py:53-60)
    w = odrv.blend_frames([np.zeros(3, np.float32)] * 8, [np.ones(3, np.float32)] * 8)
    assert np.allclose([x[0] for x in w], np.arange(8) / 7.0)


def test_library_exports_every_declared_symbol():
    from video_depth_anything_b200 import lib

    hdr = open(os.path.join(os.path.dirname(GOLD), "..", "include", "vda.h")).read()
    declared = set(re.findall(r"\b(vda_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(lib.EXPORTED), declared ^ set(lib.EXPORTED)
    if not os.path.exists(lib.LIB_PATH):
        pytest.skip("libvda_sm100a.so not built (python -c 'import __graft_entry__ as g; g.build()')")
    so = ctypes.CDLL(lib.LIB_PATH)
    for name in declared:
        assert hasattr(so, name), name
    assert lib.load().vda_version() == 1
    assert ctypes.sizeof(lib.GemmDesc) == 272


def test_product_fails_loudly_without_gpu():
    from video_depth_anything_b200 import lib

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(lib.VdaError):
        lib.require_device("cpu")

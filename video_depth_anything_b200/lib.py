"""ctypes binding of libvda_sm100a.so (include/vda.h).  The CUDA extension is the only compute path: if the
library is missing, or the device is not sm_100, every entry point raises -- there is no CPU or eager fallback."""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvda_sm100a.so")

F32, BF16 = 0, 1
MAP_SAME, MAP_DENSE, MAP_PADDED, MAP_STRIDE2, MAP_TOKENS, MAP_CONVT, MAP_S2D, MAP_CROP = range(8)
ACT_NONE, ACT_GELU, ACT_GEGLU, ACT_RELU = range(4)


class GemmDesc(C.Structure):
    """Mirror of `vda_gemm_desc` (include/vda.h) -- field order and types must match exactly."""
    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("lda", C.c_int64), ("ldb", C.c_int64), ("rowsA", C.c_int64),
        ("a_dtype", C.c_int32), ("b_dtype", C.c_int32), ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("taps", C.c_int32), ("ksplit", C.c_int32),
        ("gb", C.c_int32), ("gh", C.c_int32), ("gw", C.c_int32), ("m_padded", C.c_int32),
        ("bias", C.c_void_p), ("act", C.c_int32), ("round_bf16", C.c_int32), ("out_dtype", C.c_int32),
        ("out", C.c_void_p), ("out_relu", C.c_void_p), ("ldo", C.c_int64), ("out_map", C.c_int32),
        ("ct_k", C.c_int32), ("ct_cout", C.c_int32), ("ct_stride", C.c_int32),
        ("res", C.c_void_p), ("res2", C.c_void_p), ("ldres", C.c_int64), ("res_map", C.c_int32),
        ("resid", C.c_void_p), ("gamma", C.c_void_p), ("pos", C.c_void_p), ("ldresid", C.c_int64),
        ("resid_map", C.c_int32), ("resid_assign", C.c_int32), ("pos_rows", C.c_int32),
        ("head_w", C.c_void_p), ("head_out", C.c_void_p), ("head_b", C.c_float),
        ("use_tc", C.c_int32), ("workspace", C.c_void_p), ("workspace_bytes", C.c_int64),
    ]


_SIGS = {
    "vda_last_error": (C.c_char_p, []),
    "vda_version": (C.c_int, []),
    "vda_device_check": (C.c_int, [C.c_int, C.POINTER(C.c_int)]),
    "vda_gemm": (C.c_int, [C.POINTER(GemmDesc), C.c_void_p]),
    "vda_layernorm": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_float,
                                C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_groupnorm_tokens": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_float,
                                       C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "vda_groupnorm_tokens_noalloc": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_float,
                                               C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_groupnorm_noalloc_max_frames_x_groups": (C.c_int32, []),
    "vda_attention": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                C.c_void_p]),
    "vda_temporal_attention": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                         C.c_int32, C.c_int32, C.c_void_p]),
    "vda_rope_qk": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                              C.c_void_p]),
    "vda_bilinear": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                               C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_output_head": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int32, C.c_int32,
                                  C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_output_head_f16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int32,
                                      C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_output_head_guarded": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p,
                                          C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "vda_im2col_patch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_im2col_patch_frames": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_cls_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_pos_embed_bicubic": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_double,
                                        C.c_double, C.c_void_p]),
    "vda_preprocess_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "vda_resize_depth": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                   C.c_void_p]),
    "vda_scale_shift": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "vda_scale_shift_det_workspace_doubles": (C.c_int64, [C.c_int64]),
    "vda_scale_shift_det": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vda_scale_clamp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "vda_eval_fit": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "vda_eval_apply": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_int64, C.c_void_p,
                                 C.c_void_p]),
    "vda_scale_apply": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "vda_copy_frames": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p]),
    "vda_blend": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p]),
    "vda_copy_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "vda_host_register": (C.c_int, [C.c_void_p, C.c_int64]),
    "vda_host_unregister": (C.c_int, [C.c_void_p]),
    "vda_blend_raw": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p]),
    "vda_depth_minmax_workspace_floats": (C.c_int64, []),
    "vda_depth_minmax": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vda_depth_vis": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
}
EXPORTED = tuple(_SIGS)

_lib = None
_checked_devices = set()


class VdaError(RuntimeError):
    pass


def load():
    """dlopen the in-tree library (fails loudly when it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise VdaError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(video_depth_anything_b200/csrc/build.sh).  There is no fallback path.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def _ptr(t):
    if t is None:
        return None
    if isinstance(t, torch.Tensor):
        return t.data_ptr()
    return t


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def check(rc, what):
    if rc != 0:
        raise VdaError(f"{what} failed (rc={rc}): {load().vda_last_error().decode(errors='replace')}")


def require_device(device):
    """Raise unless `device` is an sm_100 CUDA device and the library is loadable."""
    lib = load()
    dev = torch.device(device)
    if dev.type != "cuda":
        raise VdaError(f"video_depth_anything_b200 only runs on a CUDA sm_100a device, got '{device}'")
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if idx not in _checked_devices:
        n = C.c_int(0)
        check(lib.vda_device_check(idx, C.byref(n)), "vda_device_check")
        _checked_devices.add(idx)
    return idx


def dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise VdaError(f"unsupported dtype {t.dtype}")


def gemm(desc: GemmDesc, stream=None):
    check(load().vda_gemm(C.byref(desc), stream if stream is not None else stream_ptr()), "vda_gemm")

"""The allocation-free GroupNorm entry point (`vda_groupnorm_tokens_noalloc`, include/vda.h): its statistics scratch is a
file-scope static __device__ array of the library.  Collected last on purpose (pytest runs files in name order):
a failure here cannot hide another test.

Checked through the C ABI against the existing `vda_groupnorm_tokens` (caller-provided scratch): the two run the same
kernels, so the outputs must be bit-identical; two calls in a row with different inputs and shapes show that the shared
array's left-over contents do not leak into the next call; sizes beyond the stated maximum are refused with a message;
and no device memory is allocated by the call."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(L, lib, x, w, b, frames, hw, C_, groups):
    st = L.stream_ptr()
    y_ref, y_new = torch.zeros_like(x), torch.zeros_like(x)
    scratch = torch.zeros(frames * groups * 2 * 34, dtype=torch.float64, device=x.device)
    L.check(lib.vda_groupnorm_tokens(x.data_ptr(), L.dt(x), y_ref.data_ptr(), L.dt(x), w.data_ptr(), b.data_ptr(), 1e-6,
                                     frames, hw, C_, groups, scratch.data_ptr(), st), "vda_groupnorm_tokens")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    L.check(lib.vda_groupnorm_tokens_noalloc(x.data_ptr(), L.dt(x), y_new.data_ptr(), L.dt(x), w.data_ptr(), b.data_ptr(), 1e-6,
                                             frames, hw, C_, groups, st), "vda_groupnorm_tokens_noalloc")
    torch.cuda.synchronize()
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before
    return y_ref, y_new


def test_noalloc_groupnorm_matches_existing_entry_point_twice():
    from video_depth_anything_b200 import lib as L

    lib = L.load()
    dev = torch.device("cuda", 0)
    L.require_device(dev)
    torch.cuda.set_device(dev)
    assert lib.vda_groupnorm_noalloc_max_frames_x_groups() == 2048
    g = torch.Generator(device="cpu").manual_seed(11)
    # two calls in a row, the second with other data, another shape (fewer slots of the shared array in use) and dtype
    cases = [(32, 37 * 37, 256, 32, torch.bfloat16), (3, 50, 64, 32, torch.float32), (64, 19 * 19, 1024, 32, torch.bfloat16)]
    for frames, hw, C_, groups, dt in cases:
        x = (torch.randn(frames, hw, C_, generator=g) * 1.5 + 0.3).to(device=dev, dtype=dt)
        w = (torch.rand(C_, generator=g) + 0.5).to(dev)
        b = torch.randn(C_, generator=g).to(dev)
        y_ref, y_new = _both(L, lib, x, w, b, frames, hw, C_, groups)
        assert torch.isfinite(y_new.float()).all()
        assert y_ref.float().abs().max().item() > 0.1
        assert torch.equal(y_ref, y_new), (frames, hw, C_, (y_ref.float() - y_new.float()).abs().max().item())


def test_noalloc_groupnorm_rejects_sizes_beyond_the_static_scratch():
    from video_depth_anything_b200 import lib as L

    lib = L.load()
    dev = torch.device("cuda", 0)
    L.require_device(dev)
    torch.cuda.set_device(dev)
    frames, hw, C_, groups = 65, 4, 64, 32   # 65 * 32 = 2080 > 2048
    x = torch.zeros(frames, hw, C_, device=dev)
    w, b = torch.ones(C_, device=dev), torch.zeros(C_, device=dev)
    y = torch.full_like(x, 7.0)
    rc = lib.vda_groupnorm_tokens_noalloc(x.data_ptr(), L.F32, y.data_ptr(), L.F32, w.data_ptr(), b.data_ptr(), 1e-6, frames, hw,
                                          C_, groups, L.stream_ptr())
    assert rc == -1
    msg = lib.vda_last_error().decode()
    assert "exceeds the static scratch" in msg and "2048" in msg
    torch.cuda.synchronize()
    assert bool((y == 7.0).all())  # nothing was launched

"""moca_freeinit_mix_static (no counterpart in the reference): the FreeInit frequency
mix with its spectrum scratch in a built-in device array, for callers that cannot allocate.  Checked through the C-ABI against
moca_freeinit_mix on two calls in a row (the second call finds the first one's spectrum in the shared scratch).

Collected last (file name) so that a failure here cannot hide another test under ``pytest -x``.  Tolerance: the two entry points
launch the same kernels on the same inputs, so the bar is the suite's tightest for this op (tests/test_gpu_step.py: 1e-5 rel-L2);
bit equality is what is expected and asserted.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-30))


def _call_static(lib, ctx, x, noise, lpf, y):
    from moca_video_b200._lib import ptr, stream_ptr
    T, H, W = x.shape[-3:]
    NV = x.numel() // (T * H * W)
    return lib.moca_freeinit_mix_static(ctx, ptr(x), ptr(noise), ptr(lpf), NV, T, H, W, ptr(y), stream_ptr())


def test_static_scratch_variant_matches_existing_entry_point_twice():
    from moca_video_b200 import ops
    from moca_video_b200._lib import ctx, lib
    dev = torch.device("cuda", torch.cuda.current_device())
    L, c = lib(), ctx(dev.index)
    assert L.moca_freeinit_mix_static_max_elems() == 262144
    g = torch.Generator(device=dev).manual_seed(41)
    # call 1: the FIFO shift's geometry (T = 1 volume of one 40 x 64 latent frame); call 2: a 16-frame volume that overwrites a
    # larger part of the shared scratch, then call 1's shape again on top of call 2's left-overs
    shapes = [(1, 4, 1, 40, 64), (1, 4, 16, 40, 64), (1, 4, 1, 40, 64)]
    for i, shp in enumerate(shapes):
        T, H, W = shp[-3:]
        x = torch.randn(shp, device=dev, generator=g)
        noise = torch.randn(shp, device=dev, generator=g)
        lpf = ops.freeinit_filter((T, H, W), dev, "gaussian", 1, 0.25, 0.25).to(torch.float32).reshape(T, H, W).contiguous()
        want = ops.freq_mix_3d(x, noise, lpf)                              # existing entry point (caller-provided work memory)
        y = torch.full_like(x, float("nan"))
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        rc = _call_static(L, c, x, noise, lpf, y)
        assert rc == 0, L.moca_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before                  # torch's allocator was not asked for anything
        assert torch.isfinite(y).all()
        assert _rel(y, want) < 1e-5, (i, shp)
        assert torch.equal(y, want), (i, shp)
    # the Python wrapper is the same call
    out = torch.empty_like(x)
    assert torch.equal(ops.freq_mix_3d_static(x, noise, lpf, out), want)


def test_static_scratch_variant_rejects_sizes_beyond_its_array():
    from moca_video_b200 import MocaError, ops
    from moca_video_b200._lib import ctx, lib
    dev = torch.device("cuda", torch.cuda.current_device())
    L, c = lib(), ctx(dev.index)
    shp = (2, 4, 16, 40, 64)                                                # 327680 complex values > 262144
    x = torch.zeros(shp, device=dev); y = torch.empty_like(x)
    lpf = torch.ones((16, 40, 64), device=dev)
    rc = _call_static(L, c, x, x, lpf, y)
    assert rc != 0
    msg = L.moca_last_error().decode()
    assert "262144" in msg and "moca_freeinit_mix_static" in msg
    with pytest.raises(MocaError, match="262144"):
        ops.freq_mix_3d_static(x, x, lpf, y)

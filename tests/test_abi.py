"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports exactly what include/moca_b200.h declares,
the ctypes table covers it, the weight loader expects exactly the reference's state_dict, and the product refuses to run
without a GPU (no fallback).  No compute call is made here."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "moca_b200.h")


@pytest.fixture(scope="module")
def built():
    from moca_video_b200 import build
    return build.build()


def header_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(moca_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    lib = ctypes.CDLL(built)
    names = header_functions()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/moca_b200.h but not exported"


def test_ctypes_table_matches_header(built):
    from moca_video_b200 import _lib
    assert sorted(_lib.SIGNATURES) == header_functions()
    _lib.lib()
    assert _lib.lib().moca_abi_version() == 2


def test_header_cites_reference_lines():
    src = open(HEADER).read()
    for cite in ("ddim.py:377-649", "ddim.py:274-359", "funcs.py:55-80", "funcs.py:86-99", "freeinit_utils.py:7-47",
                 "openaimodel3d.py:534-578"):
        assert cite in src, cite


def test_loader_expects_reference_state_dict(built):
    """moca_unet_create(ctx=NULL) lists the parameters the packer consumes; they must equal the reference's state_dict
    keys/sizes (arch.param_shapes is pinned to the reference listing by tests/golden/unet_full_meta.npz)."""
    import hashlib
    from moca_video_b200.arch import UNetConfig, param_shapes
    from moca_video_b200.unet import UNet3D
    u = UNet3D(_listing_only=True)
    exp = u.expected_params()
    shapes = param_shapes(UNetConfig.videocrafter2())
    assert exp == {k: int(np.prod(v)) for k, v in shapes.items()}
    g = np.load(os.path.join(ROOT, "tests", "golden", "unet_full_meta.npz"))
    listing = "\n".join(f"{k} {tuple(v)}" for k, v in sorted(shapes.items()))
    assert hashlib.sha256(listing.encode()).hexdigest() == str(g["shapes_sha256"])
    assert int(g["n_tensors"]) == len(exp) == 1484
    assert abs(u.weight_bytes() - 2 * int(g["n_params"])) / (2 * int(g["n_params"])) < 0.01     # packed 16-bit + fp32 norms/biases
    small = UNet3D(model_channels=64, context_dim=64, _listing_only=True)
    assert small.expected_params() == {k: int(np.prod(v)) for k, v in param_shapes(UNetConfig(model_channels=64, context_dim=64)).items()}


def test_unsupported_configs_fail_loudly(built):
    from moca_video_b200 import MocaError
    from moca_video_b200.unet import UNet3D
    with pytest.raises(MocaError):
        UNet3D(num_head_channels=32, _listing_only=True)          # attention kernels are built for head dim 64
    with pytest.raises(MocaError):
        UNet3D(use_linear=False, _listing_only=True)
    with pytest.raises(MocaError):
        UNet3D(model_channels=100, _listing_only=True)


def test_no_cpu_fallback(built):
    """Without a CUDA device every product entry raises instead of computing on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from moca_video_b200 import MocaError, ops
    x = torch.zeros(1, 4, 2, 4, 4)
    with pytest.raises(MocaError):
        ops.cfg_ddim_step(x, x, x, x, 1.0, np.zeros((4, 2), np.float32), [0, 0])
    with pytest.raises(MocaError):
        ops.freq_mix_3d(x, x, torch.zeros(2, 4, 4))
    with pytest.raises(MocaError):
        from moca_video_b200 import _lib
        _lib.ctx(0)
    # the ABI itself reports the failure as a status + message, never a crash
    h = ctypes.c_void_p()
    rc = _lib.lib().moca_ctx_create(0, ctypes.byref(h))
    assert rc < 0 and _lib.lib().moca_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "moca_video_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "import oracle" not in src and "from oracle" not in src, fn

"""moca_video_b200 — B200-native (sm_100a) implementation of the MoCA-Video hot path behind the reference's Python API.

All compute lives in libmoca_b200.so (moca_video_b200/csrc, C-ABI in include/moca_b200.h); this package is the thin
host-side mirror of the reference interface.  There is no CPU / PyTorch fallback: without the library, calls raise.
"""
from ._lib import LIB_PATH, MocaError  # noqa: F401


def __getattr__(name):
    import importlib
    table = {"UNet3D": ".unet", "DDIMSampler": ".sampler", "LatentDiffusionB200": ".model", "patch_reference": ".patch",
             "AutoencoderKLDecoder": ".vae", "AutoencoderKLEncoder": ".vae", "attach_vae": ".vae", "OpenCLIPTextEncoder": ".text",
             "fifo_ddim_sampling": ".fifo", "prepare_latents": ".fifo", "shift_latents": ".fifo"}
    if name in table:
        return getattr(importlib.import_module(table[name], __name__), name)
    if name in ("ops", "fifo", "sampler", "schedule", "unet", "model", "patch", "arch", "build", "vae", "text", "driver"):
        return importlib.import_module("." + name, __name__)
    raise AttributeError(name)

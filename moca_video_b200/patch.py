"""``patch_reference()``: install the B200 path behind the reference's own Python surface.

The reference keeps every class, constructor, ``state_dict`` key and method signature; only the bodies of the methods on
the hot path are rebound to call libmoca_b200.so — the same module-level rebinding the reference itself uses to switch
CrossAttention to xformers (lvdm/modules/attention.py:73-74).  Call after ``sys.path`` contains the reference tree:

    import moca_video_b200 as mb; mb.patch_reference()

Rebound:
  lvdm.modules.networks.openaimodel3d.UNetModel.forward         -> UNet3D.forward (weights packed on first call)
  lvdm.models.samplers.ddim.DDIMSampler.unet / fifo_onestep / ddim_step / p_sample_ddim / make_schedule
  utils.freeinit_utils.freq_mix_3d / get_freq_filter
  scripts.evaluation.funcs.prepare_latents / shift_latents / fifo_ddim_sampling
"""
import importlib

import torch

from . import fifo, ops
from .sampler import DDIMSampler as _B200Sampler
from .unet import UNet3D


def _unet_forward(self, x, timesteps, context=None, features_adapter=None, fps=16, **kwargs):
    """Replacement body of UNetModel.forward (openaimodel3d.py:534-578)."""
    eng = getattr(self, "_moca_engine", None)
    if eng is None:
        mult = tuple(self.channel_mult) if hasattr(self, "channel_mult") else (1, 2, 4, 4)
        eng = UNet3D(self.in_channels, self.out_channels, self.model_channels, tuple(self.attention_resolutions),
                     self.num_res_blocks, mult, 64, _context_dim(self), device=x.device)
        eng.load_state_dict(self.state_dict())
        object.__setattr__(self, "_moca_engine", eng)
    return eng.forward(x, timesteps, context=context, fps=fps)     # [b] or per-frame [b*t] timesteps (is_fifo, :535)


def _context_dim(unet):
    for m in unet.modules():
        if m.__class__.__name__ == "SpatialTransformer":
            return m.transformer_blocks[0].attn2.to_k.in_features
    return 1024


def _get_freq_filter(shape, device, filter_type, n, d_s, d_t):
    """utils/freeinit_utils.py:51-70; returns a [T,H,W] tensor that broadcasts like the reference's mask."""
    return ops.freeinit_filter(shape, device, filter_type, n, d_s, d_t)


def _freq_mix_3d(x, noise, LPF):
    """utils/freeinit_utils.py:7-47 (the reference squeezes dim 0 and returns [C,T,H,W] for b=1)."""
    y = ops.freq_mix_3d(x, noise, LPF)
    return y.squeeze(0)


def patch_reference():
    mods = {}
    for name in ("lvdm.modules.networks.openaimodel3d", "lvdm.models.samplers.ddim", "utils.freeinit_utils",
                 "scripts.evaluation.funcs"):
        try:
            mods[name] = importlib.import_module(name)
        except Exception as e:  # noqa: BLE001  (a caller may only use part of the surface)
            mods[name] = None
            last = e
    if all(m is None for m in mods.values()):
        raise ImportError(f"patch_reference: the reference tree is not importable ({last})")
    m = mods["lvdm.modules.networks.openaimodel3d"]
    if m is not None:
        m.UNetModel.forward = _unet_forward
    m = mods["lvdm.models.samplers.ddim"]
    if m is not None:
        S = m.DDIMSampler
        for meth in ("make_schedule", "unet", "_eps_pair", "fifo_onestep", "ddim_step", "_step", "p_sample_ddim"):
            setattr(S, meth, getattr(_B200Sampler, meth))
        S.parity_quirks = True
        S.noise_fn = staticmethod(lambda shape, device, generator=None: torch.randn(shape, device=device))
    m = mods["utils.freeinit_utils"]
    if m is not None:
        m.get_freq_filter = _get_freq_filter
        m.freq_mix_3d = _freq_mix_3d
    m = mods["scripts.evaluation.funcs"]
    if m is not None:
        m.prepare_latents = fifo.prepare_latents
        m.shift_latents = fifo.shift_latents
        m.get_freq_filter = _get_freq_filter
        m.freq_mix_3d = _freq_mix_3d
    return {k: v is not None for k, v in mods.items()}

"""``patch_reference()``: install the B200 path behind the reference's own Python surface.

The reference keeps every class, constructor, ``state_dict`` key and method signature; only the bodies of the methods on
the hot path are rebound to call libmoca_b200.so — the same module-level rebinding the reference itself uses to switch
CrossAttention to xformers (lvdm/modules/attention.py:73-74).  Call after ``sys.path`` contains the reference tree:

    import moca_video_b200 as mb; mb.patch_reference()

Rebound (``REBOUND`` lists them; tests/test_patch.py checks each against the reference's own signature):
  lvdm.modules.networks.openaimodel3d.UNetModel.forward          -> UNet3D.forward (weights packed on first call)
  lvdm.modules.networks.openaimodel3d.ResBlock.forward           -> UNet3D.layer_forward (moca_unet_layer_forward), incl. its
                                                                    TemporalConvBlock
  lvdm.modules.attention.SpatialTransformer.forward / TemporalTransformer.forward -> UNet3D.layer_forward
      (a module reaches the engine of the UNetModel it belongs to: ``attach(unet)``, done by the first UNetModel.forward)
  lvdm.models.samplers.ddim.DDIMSampler.make_schedule / unet / fifo_onestep / ddim_step / p_sample_ddim
  utils.freeinit_utils.freq_mix_3d / get_freq_filter
  scripts.evaluation.funcs.prepare_latents / shift_latents / fifo_ddim_sampling (+ its imported FreeInit names)
  lvdm.models.autoencoder.AutoencoderKL.decode / .encode         -> AutoencoderKLDecoder.decode (moca_vae_decode) /
                                                                    AutoencoderKLEncoder.encode_moments (moca_vae_encode); SURVEY N1
  lvdm.models.ddpm3d.LatentDiffusion.decode_first_stage_2DAE     -> the same decoder, frames grouped per call instead of one
                                                                    ``decode`` per frame (ddpm3d.py:555-562)
      (both need pytorch_lightning to import; where it is absent they are skipped and reported False)
  lvdm.modules.encoders.condition.FrozenOpenCLIPEmbedder.encode_with_transformer -> OpenCLIPTextEncoder (moca_text_encode);
      tokenisation (``forward`` -> open_clip.tokenize) stays the reference's
The rebound ``fifo_ddim_sampling`` calls ``model.decode_first_stage_2DAE`` on every popped frame, exactly where the reference
does (funcs.py:359-365).  
"""
import importlib
import os

import torch

from . import fifo, ops
from ._lib import MocaError
from .sampler import DDIMSampler as _B200Sampler
from .unet import UNet3D
from .text import attach_text
from .vae import DiagonalGaussian, attach_vae, attach_vae_encoder

REBOUND = {
    "lvdm.modules.networks.openaimodel3d": ["UNetModel.forward", "ResBlock.forward"],
    "lvdm.modules.attention": ["SpatialTransformer.forward", "TemporalTransformer.forward"],
    "lvdm.models.samplers.ddim": ["DDIMSampler.make_schedule", "DDIMSampler.unet", "DDIMSampler.fifo_onestep", "DDIMSampler.ddim_step",
                                  "DDIMSampler.p_sample_ddim"],
    "lvdm.models.autoencoder": ["AutoencoderKL.decode", "AutoencoderKL.encode"],
    "lvdm.models.ddpm3d": ["LatentDiffusion.decode_first_stage_2DAE"],
    "lvdm.modules.encoders.condition": ["FrozenOpenCLIPEmbedder.encode_with_transformer"],
    "utils.freeinit_utils": ["freq_mix_3d", "get_freq_filter"],
    "scripts.evaluation.funcs": ["prepare_latents", "shift_latents", "fifo_ddim_sampling", "freq_mix_3d", "get_freq_filter"],
}

# constructor options of UNetModel that select code the library implements; anything else must fail loudly (UNet3D checks)
_UNET_FLAGS = ("use_linear", "temporal_conv", "temporal_attention", "temporal_selfatt_only", "use_relative_position",
               "use_causal_attention", "addition_attention", "fps_cond", "image_cross_attention", "use_scale_shift_norm",
               "resblock_updown", "conv_resample", "dims")


def _head_channels(unet):
    for m in unet.modules():
        if m.__class__.__name__ == "SpatialTransformer":
            a = m.transformer_blocks[0].attn1
            return a.to_q.out_features // a.heads
    return 64


def _context_dim(unet):
    for m in unet.modules():
        if m.__class__.__name__ == "SpatialTransformer":
            return m.transformer_blocks[0].attn2.to_k.in_features
    return 1024


def _init_attn_heads(unet):
    ia = getattr(unet, "init_attn", None)
    if ia is None:
        return 8
    for m in ia.modules():
        if m.__class__.__name__ == "CrossAttention":
            return m.heads
    return 8


_MODULE_KINDS = ("ResBlock", "SpatialTransformer", "TemporalTransformer")


def attach(unet, device=None):
    """Builds (once) the library engine of a reference ``UNetModel`` — weights packed from its ``state_dict`` — and tells every
    ResBlock / SpatialTransformer / TemporalTransformer inside which engine and which layer name it is, so that the rebound module
    ``forward``s reach the same packed weights."""
    eng = getattr(unet, "_moca_engine", None)
    if eng is None:
        dev = device if device is not None else next(unet.parameters()).device
        mult = tuple(unet.channel_mult) if hasattr(unet, "channel_mult") else (1, 2, 4, 4)
        flags = {k: getattr(unet, k) for k in _UNET_FLAGS if hasattr(unet, k)}      # unsupported options raise in UNet3D
        eng = UNet3D(unet.in_channels, unet.out_channels, unet.model_channels, tuple(unet.attention_resolutions),
                     unet.num_res_blocks, mult, _head_channels(unet), _context_dim(unet), _init_attn_heads(unet), device=dev, **flags)
        eng.load_state_dict(unet.state_dict())
        object.__setattr__(unet, "_moca_engine", eng)
        for name, mod in unet.named_modules():
            if mod.__class__.__name__ in _MODULE_KINDS:
                object.__setattr__(mod, "_moca_engine", eng)
                object.__setattr__(mod, "_moca_name", name)
    return eng


def _unet_forward(self, x, timesteps, context=None, features_adapter=None, fps=16, ctx_len=None, **kwargs):
    """Replacement body of UNetModel.forward (openaimodel3d.py:534-578).  ``ctx_len`` is the one addition: true context
    lengths when the caller batches samples with different prompt lengths (the patched sampler's b=2 CFG launch)."""
    eng = attach(self, x.device)
    return eng.forward(x, timesteps, context=context, features_adapter=features_adapter, fps=fps, ctx_len=ctx_len)


def _module_engine(mod, what):
    eng = getattr(mod, "_moca_engine", None)
    if eng is None:
        raise MocaError(f"{what}: this module is not attached to a library engine; call moca_video_b200.patch.attach(unet) "
                        "on the UNetModel it belongs to (the first patched UNetModel.forward does it)")
    return eng, mod._moca_name


def _resblock_forward(self, x, emb, batch_size=None):
    """ResBlock.forward (openaimodel3d.py:195-234) with its TemporalConvBlock (:269-276): x [(b t), c, h, w], emb [(b t), 4 mc]."""
    eng, name = _module_engine(self, "ResBlock.forward")
    if not batch_size:
        raise MocaError("ResBlock.forward: batch_size is required (the library always runs the block's temporal convolution, "
                        "which the reference skips without it, openaimodel3d.py:230)")
    n, c, h, w = x.shape
    b, t = int(batch_size), n // int(batch_size)
    y = eng.layer_forward(name, x.reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4), emb=emb)
    return y.permute(0, 2, 1, 3, 4).reshape(n, y.shape[1], h, w)


def _spatial_tf_forward(self, x, context=None):
    """SpatialTransformer.forward (attention.py:262-278): x [(b t), c, h, w], context [(b t), L, d] (the same text for the frames
    of a sample).  Frames are independent here, so any grouping into (b, t) with one context per group is equivalent."""
    eng, name = _module_engine(self, "SpatialTransformer.forward")
    if context is None:
        raise MocaError("SpatialTransformer.forward: context is required (cross-attention on the text embedding)")
    n, c, h, w = x.shape
    t = 16 if n % 16 == 0 else 1
    ctx = context.reshape(n // t, t, *context.shape[1:])
    if t > 1 and not bool((ctx == ctx[:, :1]).all()):
        t = 1
        ctx = context.reshape(n, 1, *context.shape[1:])
    b = n // t
    if b > 16:
        raise MocaError(f"SpatialTransformer.forward: {n} frames with per-frame contexts exceed one launch (16 groups)")
    y = eng.layer_forward(name, x.reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4), context=ctx[:, 0])
    return y.permute(0, 2, 1, 3, 4).reshape(n, c, h, w)


def _temporal_tf_forward(self, x, context=None):
    """TemporalTransformer.forward (attention.py:331-373), temporal_selfatt_only: x [b, c, t, h, w]; context is unused there too."""
    eng, name = _module_engine(self, "TemporalTransformer.forward")
    return eng.layer_forward(name, x)


def _ae_decode(self, z, **kwargs):
    """Replacement body of AutoencoderKL.decode (autoencoder.py:104-107): z [n, zc, h, w], already divided by scale_factor."""
    return attach_vae(self, device=z.device).decode(z).to(z.dtype)


def _clip_encode_with_transformer(self, text):
    """Replacement body of FrozenOpenCLIPEmbedder.encode_with_transformer (condition.py:212-219): token ids [b, 77] -> [b, 77, 1024].
    ``forward`` (tokenisation with open_clip.tokenize, condition.py:205-210) stays the reference's."""
    return attach_text(self, device=text.device if text.is_cuda else None).encode_with_transformer(text)


def _ae_encode(self, x, **kwargs):
    """Replacement body of AutoencoderKL.encode (autoencoder.py:97-102): the posterior over z for images x [n, 3, H, W]; the
    returned object has the reference distribution's ``sample`` / ``mode`` / ``parameters`` (lvdm/distributions.py:24-43), which is
    what get_first_stage_encoding (ddpm3d.py:458-465) consumes; that function tests isinstance, so the reference class is used when it
    imports."""
    mom = attach_vae_encoder(self, device=x.device).encode_moments(x).to(x.dtype)
    try:
        from lvdm.distributions import DiagonalGaussianDistribution
        return DiagonalGaussianDistribution(mom)
    except ImportError:
        return DiagonalGaussian(mom)


def _decode_first_stage_2DAE(self, z, **kwargs):
    """Replacement body of LatentDiffusion.decode_first_stage_2DAE (ddpm3d.py:555-562): z [b, zc, t, h, w]."""
    eng = attach_vae(self.first_stage_model, device=z.device)
    eng.scale_factor = float(self.scale_factor)
    return eng.decode_first_stage_2DAE(z).to(z.dtype)


def _get_freq_filter(shape, device, filter_type, n, d_s, d_t):
    """utils/freeinit_utils.py:51-70; returns a [T,H,W] tensor that broadcasts like the reference's mask."""
    return ops.freeinit_filter(shape, device, filter_type, n, d_s, d_t)


def _freq_mix_3d(x, noise, LPF):
    """utils/freeinit_utils.py:7-47 (the reference squeezes dim 0 and returns [C,T,H,W] for b=1)."""
    y = ops.freq_mix_3d(x, noise, LPF)
    return y.squeeze(0)


def _make_fifo_ddim_sampling(funcs_mod):
    def fifo_ddim_sampling(args, model, conditioning, noise_shape, ddim_sampler, cfg_scale=1.0, output_dir=None, latents_dir=None,
                           save_frames=False, conditioned_image=None, targets=None, gamma=0.5, use_self_attention=False,
                           davis_data=None, anchor_frame=None, **kwargs):
        """scripts/evaluation/funcs.py:243-371 with the latent loop in the library (windows sharded over the ranks of the
        default process group when there is one); every popped latent frame goes through the reference's own VAE decoder
        and ``tensor2image`` (funcs.py:359-365).  Returns the list of PIL frames like the reference."""
        if davis_data is not None:
            raise NotImplementedError("fifo_ddim_sampling: the DAVIS-video branch needs the VAE encoder in shift_latents (SURVEY N1)")
        frames = []
        fifo_dir = None
        if getattr(args, "save_frames", False):                       # funcs.py:283-285
            fifo_dir = os.path.join(output_dir, "fifo")
            os.makedirs(fifo_dir, exist_ok=True)

        def on_frame(i, z):
            image = funcs_mod.tensor2image(model.decode_first_stage_2DAE(z))
            if save_frames and fifo_dir is not None:
                image.save(os.path.join(fifo_dir, f"{i}.png"))
            frames.append(image)
        fifo.fifo_ddim_sampling(args, model, conditioning, noise_shape, ddim_sampler, cfg_scale=cfg_scale, output_dir=output_dir,
                                latents_dir=latents_dir, save_frames=save_frames, conditioned_image=conditioned_image,
                                targets=targets, gamma=gamma, use_self_attention=use_self_attention, on_frame=on_frame, **kwargs)
        return frames
    return fifo_ddim_sampling


def patch_reference():
    mods = {}
    last = None
    for name in REBOUND:
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
        m.ResBlock.forward = _resblock_forward
    m = mods["lvdm.modules.attention"]
    if m is not None:
        m.SpatialTransformer.forward = _spatial_tf_forward
        m.TemporalTransformer.forward = _temporal_tf_forward
    # (under a stubbed pytorch_lightning the two classes below are not real classes: leave them alone)
    m = mods["lvdm.models.autoencoder"]
    if m is not None and isinstance(getattr(m, "AutoencoderKL", None), type):
        m.AutoencoderKL.decode = _ae_decode
        m.AutoencoderKL.encode = _ae_encode
    else:
        mods["lvdm.models.autoencoder"] = None
    m = mods["lvdm.models.ddpm3d"]
    if m is not None and isinstance(getattr(m, "LatentDiffusion", None), type):
        m.LatentDiffusion.decode_first_stage_2DAE = _decode_first_stage_2DAE
    else:
        mods["lvdm.models.ddpm3d"] = None
    m = mods["lvdm.modules.encoders.condition"]
    if m is not None and isinstance(getattr(m, "FrozenOpenCLIPEmbedder", None), type):
        m.FrozenOpenCLIPEmbedder.encode_with_transformer = _clip_encode_with_transformer
    else:
        mods["lvdm.modules.encoders.condition"] = None
    m = mods["lvdm.models.samplers.ddim"]
    if m is not None:
        S = m.DDIMSampler
        for meth in ("make_schedule", "unet", "_eps_pair", "fifo_onestep", "ddim_step", "_step", "_step_sam", "p_sample_ddim"):
            setattr(S, meth, getattr(_B200Sampler, meth))
        S.parity_quirks = True
        S.segment_fn = None
        S.noise_fn = staticmethod(lambda shape, device, generator=None: torch.randn(shape, device=device))
    m = mods["utils.freeinit_utils"]
    if m is not None:
        m.get_freq_filter = _get_freq_filter
        m.freq_mix_3d = _freq_mix_3d
    m = mods["scripts.evaluation.funcs"]
    if m is not None:
        m.prepare_latents = fifo.prepare_latents
        m.shift_latents = fifo.shift_latents
        m.fifo_ddim_sampling = _make_fifo_ddim_sampling(m)
        m.get_freq_filter = _get_freq_filter
        m.freq_mix_3d = _freq_mix_3d
    return {k: v is not None for k, v in mods.items()}

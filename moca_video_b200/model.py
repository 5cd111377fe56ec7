"""Host-side stand-in for the part of ``LatentDiffusion`` the sampler touches (lvdm/models/ddpm3d.py).

Carries the DDPM schedule buffers (register_schedule, ddpm3d.py:113-141), ``use_scale`` / ``scale_arr`` (:362-376) and
``apply_model`` (:512-527 -> DiffusionWrapper.forward :702-712: concatenate ``c_crossattn`` along tokens, pass ``fps``).
``first_stage`` / ``first_stage_encoder`` (optional): ``AutoencoderKLDecoder`` / ``AutoencoderKLEncoder`` (vae.py) behind
``decode_first_stage_2DAE`` (:555-562) and ``encode_first_stage_2DAE`` (:485-490).
``cond_stage`` (optional): an ``OpenCLIPTextEncoder`` (text.py) behind ``get_learned_conditioning`` (:445-456).
"""
import numpy as np
import torch

from .schedule import cumprod_alphas_f32, linear_betas, scale_array_f32
from .unet import UNet3D


class LatentDiffusionB200:
    def __init__(self, unet: UNet3D, timesteps=1000, linear_start=0.00085, linear_end=0.012, use_scale=True, fps=10,
                 first_stage=None, scale_factor=0.18215, first_stage_encoder=None, cond_stage=None):
        self.unet = unet
        self.first_stage_model = first_stage
        self.scale_factor = float(scale_factor)
        self.first_stage_encoder = first_stage_encoder
        self.cond_stage_model = cond_stage
        for part in (first_stage, first_stage_encoder):
            if part is not None:
                part.scale_factor = self.scale_factor
        self.device = unet.device
        self.num_timesteps = timesteps
        betas = linear_betas(timesteps, linear_start, linear_end)
        ac = cumprod_alphas_f32(timesteps, linear_start, linear_end)
        self.betas = torch.tensor(betas, dtype=torch.float32, device=self.device)
        self.alphas_cumprod = torch.tensor(ac, dtype=torch.float32)
        self.alphas_cumprod_prev = torch.tensor(np.append(np.float32(1.0), ac[:-1]), dtype=torch.float32)
        self.use_scale = use_scale
        self.scale_arr = torch.tensor(scale_array_f32(), dtype=torch.float32)
        self.default_fps = fps

    def decode_first_stage_2DAE(self, z, **kwargs):
        """ddpm3d.py:555-562: z [b, zc, t, h, w] sampler latents -> frames [b, 3, t, 8h, 8w] in [-1, 1]."""
        if self.first_stage_model is None:
            from ._lib import MocaError
            raise MocaError("LatentDiffusionB200.decode_first_stage_2DAE: constructed without first_stage (an AutoencoderKLDecoder)")
        return self.first_stage_model.decode_first_stage_2DAE(z)

    def get_learned_conditioning(self, c):
        """ddpm3d.py:445-456: list of prompts (or token ids [b, 77]) -> text embedding [b, 77, 1024]."""
        if self.cond_stage_model is None:
            from ._lib import MocaError
            raise MocaError("LatentDiffusionB200.get_learned_conditioning: constructed without cond_stage (an OpenCLIPTextEncoder)")
        if isinstance(c, torch.Tensor):
            return self.cond_stage_model.encode_with_transformer(c)
        return self.cond_stage_model.encode(c)

    def encode_first_stage_2DAE(self, x, noise=None):
        """ddpm3d.py:485-490: x [b, 3, t, H, W] frames in [-1, 1] -> scale_factor * sample(posterior) [b, zc, t, H/8, W/8]."""
        if self.first_stage_encoder is None:
            from ._lib import MocaError
            raise MocaError("LatentDiffusionB200.encode_first_stage_2DAE: constructed without first_stage_encoder (an AutoencoderKLEncoder)")
        return self.first_stage_encoder.encode_first_stage_2DAE(x, noise=noise)

    @staticmethod
    def _context(c):
        """DiffusionWrapper.forward 'crossattn' branch (ddpm3d.py:710-712): cat the c_crossattn list along tokens."""
        if isinstance(c, dict):
            ctx = torch.cat(c["c_crossattn"], dim=1) if isinstance(c["c_crossattn"], (list, tuple)) else c["c_crossattn"]
            return ctx, c.get("fps", None)
        return c, None

    def apply_model(self, x_noisy, t, cond, **kwargs):
        ctx, fps = self._context(cond)
        return self.unet(x_noisy, t, context=ctx, fps=self.default_fps if fps is None else fps, **kwargs)

    def apply_model_cfg(self, x_noisy, t, cond, uncond, **kwargs):
        """The two apply_model calls of classifier-free guidance (ddim.py:368-370) as one b=2 UNet launch per sample.
        Contexts of different length (154 vs 77 tokens) are padded and carry their true length."""
        c_ctx, fps = self._context(cond)
        u_ctx, _ = self._context(uncond)
        b = x_noisy.shape[0]
        Lc, Lu = c_ctx.shape[1], u_ctx.shape[1]
        L = max(Lc, Lu)
        ctx = torch.zeros((2 * b, L, c_ctx.shape[2]), device=self.device, dtype=torch.float32)
        ctx[:b, :Lc] = c_ctx
        ctx[b:, :Lu] = u_ctx
        x2 = torch.cat([x_noisy, x_noisy], dim=0)
        ts = list(t.tolist() if isinstance(t, torch.Tensor) else t)
        fps = self.default_fps if fps is None else fps
        if isinstance(fps, torch.Tensor):
            fps = fps.reshape(-1)
            if fps.numel() == 1:
                fps = fps.repeat(2 * b)
            elif fps.numel() == b:
                fps = torch.cat([fps, fps])
            else:
                raise ValueError(f"apply_model_cfg: fps has {fps.numel()} entries for batch {b} (1 or {b} expected)")
        e = self.unet(x2, ts + ts, context=ctx, fps=fps, ctx_len=[Lc] * b + [Lu] * b, **kwargs)
        return e[:b], e[b:]

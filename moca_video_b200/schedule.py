"""Host-side coefficient tables of the sampler (no compute on tensors: 50-64 scalars per table).

Restates, with the reference's mixed precisions, what these reference functions produce:
  make_beta_schedule("linear")            lvdm/models/utils_diffusion.py:31-36
  DDPM.register_schedule                  lvdm/models/ddpm3d.py:113-141   (alphas_cumprod as a float32 buffer)
  LatentDiffusion.scale_arr               lvdm/models/ddpm3d.py:362-376
  make_ddim_timesteps("uniform")          lvdm/models/utils_diffusion.py:56-66
  make_ddim_sampling_parameters           lvdm/models/utils_diffusion.py:81-93
  DDIMSampler.make_schedule               lvdm/models/samplers/ddim.py:62-106
The tables are the *inputs* of the CUDA step kernels (``float[4][F]`` blocks, include/moca_b200.h).
"""
from dataclasses import dataclass

import numpy as np


def linear_betas(num=1000, start=0.00085, end=0.012):
    return np.square(np.linspace(np.sqrt(start), np.sqrt(end), num, dtype=np.float64))


def cumprod_alphas_f32(num=1000, start=0.00085, end=0.012):
    return np.cumprod(1.0 - linear_betas(num, start, end)).astype(np.float32)


def scale_array_f32(scale_a=1.0, scale_b=0.7, ramp=400, tail=1000):
    ramp_part = np.linspace(scale_a, scale_b, ramp)
    return np.concatenate([ramp_part, np.full(tail, scale_b)]).astype(np.float32)


def ddim_timesteps(num_ddim, num_ddpm=1000):
    return np.round(np.linspace(0, num_ddpm - 1, num_ddim)).astype(np.int64)


@dataclass
class DDIMTables:
    timesteps: np.ndarray             # int64 [S]
    alphas: np.ndarray                # float32 [S]
    alphas_prev: np.ndarray           # float32 [S] (float64 ndarray in the reference, read into float32 tensors)
    sigmas: np.ndarray                # float32 [S]
    sqrt_one_minus_alphas: np.ndarray  # float32 [S]
    scale: np.ndarray                 # float32 [S]
    scale_prev: np.ndarray            # float32 [S]

    def coef(self, indices):
        """float32 [4, F] block (A, A_prev, sigma, sqrt(1-A)) for the frames' table indices."""
        ix = np.asarray(indices, dtype=np.int64)
        return np.ascontiguousarray(np.stack([self.alphas[ix], self.alphas_prev[ix], self.sigmas[ix],
                                              self.sqrt_one_minus_alphas[ix]]).astype(np.float32))


def make_ddim_tables(num_ddim, eta, alphas_cumprod=None, scale_arr=None, num_ddpm=1000):
    ac = cumprod_alphas_f32(num_ddpm) if alphas_cumprod is None else np.asarray(alphas_cumprod, dtype=np.float32)
    sc = scale_array_f32() if scale_arr is None else np.asarray(scale_arr, dtype=np.float32)
    steps = ddim_timesteps(num_ddim, num_ddpm)
    a = ac[steps]                                                        # float32, as the reference's tensor indexing
    a_prev64 = np.concatenate([[np.float64(ac[0])], ac[steps[:-1]].astype(np.float64)])
    # sigma = eta * sqrt((1 - a_prev) / (1 - a) * (1 - a / a_prev)); the reference divides a float64 ndarray by a
    # float32 tensor, i.e. multiplies by the float32 reciprocal of (1 - a)
    inv_one_minus_a = (np.float32(1.0) / (np.float32(1.0) - a)).astype(np.float64)
    sig = eta * np.sqrt(inv_one_minus_a * (1.0 - a_prev64) * (1.0 - a.astype(np.float64) / a_prev64))
    s1m = np.sqrt(np.float32(1.0) - a).astype(np.float32)
    scale = sc[steps]
    scale_prev = np.concatenate([[sc[0]], sc[steps[:-1]]]).astype(np.float32)
    return DDIMTables(steps, a.astype(np.float32), a_prev64.astype(np.float32), sig.astype(np.float32), s1m,
                      scale.astype(np.float32), scale_prev)


def fifo_queue_tables(num_steps=64, video_length=16, lookahead=True, num_ddpm=1000):
    """Queue-slot timesteps and table indices (scripts/evaluation/funcs.py:290-294)."""
    ts = ddim_timesteps(num_steps, num_ddpm)
    idx = np.arange(num_steps)
    if lookahead:
        half = video_length // 2
        ts = np.concatenate([np.repeat(ts[:1], half), ts])
        idx = np.concatenate([np.zeros(half, dtype=idx.dtype), idx])
    return ts, idx


def reference_mask_index(num_frames, height):
    """Mask frame each latent frame reads in the reference AS SHIPPED: the frame loop variable is overwritten by the
    visualisation loops (ddim.py:405,477,502,533), so frames >= 1 read mask ceil(H/4)-1."""
    j = -(-height // 4) - 1
    return [0] + [j] * (num_frames - 1)

"""Import the unmodified reference (``/root/reference``) in a GPU-less container.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Works only where the reference tree
is mounted (the authoring container); nothing that runs on the GPU box may call this.

Recipe follows SURVEY.md Appendix C: third-party modules that are absent here (``sam2``,
``matplotlib``, ``decord`` ...) are replaced by ``MagicMock`` *before* import; the sampler is
sub-classed so ``register_buffer`` does not force tensors to "cuda" (ddim.py:53-60) and is
constructed with ``use_self_attention=True`` so it does not try to load SAM-2 (ddim.py:49-50).
"""
import os
import sys
import types
from unittest.mock import MagicMock

import numpy as np
import torch

REF = os.environ.get("MOCA_REFERENCE", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "lvdm"))


def _stub(name):
    if name not in sys.modules:
        sys.modules[name] = MagicMock()


def setup():
    if not available():
        raise RuntimeError("reference tree not present: " + REF)
    sys.dont_write_bytecode = True
    if REF not in sys.path:
        sys.path.insert(0, REF)
    try:
        __import__("pytorch_lightning")
    except Exception:
        # AutoencoderKL / DDPM subclass pl.LightningModule (autoencoder.py:13, ddpm3d.py:47): a plain nn.Module base keeps them
        # real classes (constructors, state_dict, decode), which is all the parity fixtures and the patch tests use
        pl = MagicMock()
        pl.LightningModule = torch.nn.Module
        sys.modules["pytorch_lightning"] = pl
    for m in ("sam2", "sam2.build_sam", "sam2.sam2_image_predictor", "matplotlib", "matplotlib.pyplot",
              "matplotlib.cm", "decord", "imageio", "cv2", "omegaconf", "kornia", "torchvision", "torchvision.utils",
              "open_clip"):
        try:
            __import__(m)
        except Exception:
            _stub(m)


def unet_params():
    import yaml
    with open(os.path.join(REF, "configs/inference_t2v_512_v2.0.yaml")) as f:
        cfg = yaml.safe_load(f)
    p = dict(cfg["model"]["params"]["unet_config"]["params"])
    p["use_checkpoint"] = False
    return p


def build_unet(params=None, seed=0, rerandomize=True):
    """Reference ``UNetModel`` (openaimodel3d.py:279) with seeded weights; zero-initialised tensors
    re-drawn N(0, 0.02^2) so the output is not trivially zero."""
    setup()
    from lvdm.modules.networks.openaimodel3d import UNetModel
    p = params or unet_params()
    torch.manual_seed(seed)
    net = UNetModel(**p).eval()
    if rerandomize:
        rerandomize_zero_params(net, seed + 1)
    return net


def rerandomize_zero_params(net, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for _, prm in sorted(net.named_parameters()):
            if prm.numel() > 0 and float(prm.abs().max()) == 0.0:
                prm.copy_(torch.randn(prm.shape, generator=g) * 0.02)


class ShimModel:
    """Stands in for ``LatentDiffusion`` (ddpm3d.py:113-165, 362-376, 512-527): schedule buffers + apply_model."""

    def __init__(self, eps_fn=None, use_scale=True):
        setup()
        from lvdm.models.utils_diffusion import make_beta_schedule
        betas = make_beta_schedule("linear", 1000, linear_start=0.00085, linear_end=0.012)
        alphas = 1.0 - betas
        ac = np.cumprod(alphas, axis=0)
        acp = np.append(1.0, ac[:-1])
        self.num_timesteps = 1000
        self.betas = torch.tensor(betas, dtype=torch.float32)
        self.alphas_cumprod = torch.tensor(ac, dtype=torch.float32)
        self.alphas_cumprod_prev = torch.tensor(acp, dtype=torch.float32)
        self.use_scale = use_scale
        sa = np.concatenate((np.linspace(1.0, 0.7, 400), np.full(1000, 0.7)))
        self.scale_arr = torch.tensor(sa, dtype=torch.float32)
        self.device = torch.device("cpu")
        self.eps_fn = eps_fn
        self.calls = []

    def apply_model(self, x, t, c, **kw):
        self.calls.append(dict(kw))
        return self.eps_fn(x, t, c)


def make_sampler(model, steps, eta):
    setup()
    from lvdm.models.samplers.ddim import DDIMSampler

    class CPUSampler(DDIMSampler):
        def register_buffer(self, name, attr):
            setattr(self, name, attr)

    s = CPUSampler(model, use_self_attention=True)
    s.make_schedule(steps, ddim_eta=eta, verbose=False)
    return s

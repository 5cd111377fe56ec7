"""patch_reference(): every listed symbol of the reference is rebound, and every replacement accepts the
reference's own call.  Needs the reference tree (authoring container); CPU only — the calls must reach the device boundary
and stop there with MocaError (there is no CPU fallback), never with a TypeError."""
import inspect
import os
import sys
import types

import numpy as np
import pytest
import torch

from oracle import refimport

pytestmark = pytest.mark.skipif(not refimport.available(), reason="reference tree not mounted")

ORIGINAL_AT = {                                   # reference definitions the replacements stand in for (file:line)
    "lvdm.modules.networks.openaimodel3d:UNetModel.forward": "openaimodel3d.py:534",
    "lvdm.modules.networks.openaimodel3d:ResBlock.forward": "openaimodel3d.py:195",
    "lvdm.modules.attention:SpatialTransformer.forward": "attention.py:262",
    "lvdm.modules.attention:TemporalTransformer.forward": "attention.py:331",
    "lvdm.models.samplers.ddim:DDIMSampler.make_schedule": "ddim.py:62",
    "lvdm.models.samplers.ddim:DDIMSampler.unet": "ddim.py:362",
    "lvdm.models.samplers.ddim:DDIMSampler.fifo_onestep": "ddim.py:255",
    "lvdm.models.samplers.ddim:DDIMSampler.ddim_step": "ddim.py:377",
    "lvdm.models.samplers.ddim:DDIMSampler.p_sample_ddim": "ddim.py:274",
    "lvdm.models.autoencoder:AutoencoderKL.decode": "autoencoder.py:104",
    "lvdm.models.autoencoder:AutoencoderKL.encode": "autoencoder.py:97",
    "lvdm.models.ddpm3d:LatentDiffusion.decode_first_stage_2DAE": "ddpm3d.py:556",
    "lvdm.modules.encoders.condition:FrozenOpenCLIPEmbedder.encode_with_transformer": "condition.py:212",
    "utils.freeinit_utils:freq_mix_3d": "freeinit_utils.py:7",
    "utils.freeinit_utils:get_freq_filter": "freeinit_utils.py:51",
    "scripts.evaluation.funcs:prepare_latents": "funcs.py:21",
    "scripts.evaluation.funcs:shift_latents": "funcs.py:86",
    "scripts.evaluation.funcs:fifo_ddim_sampling": "funcs.py:243",
}


def _resolve(mod, dotted):
    obj = mod
    for part in dotted.split("."):
        obj = getattr(obj, part)
    return obj


@pytest.fixture(scope="module")
def patched():
    """(originals, modules) — originals captured before patching."""
    import importlib
    refimport.setup()
    import moca_video_b200.patch as P
    mods = {name: importlib.import_module(name) for name in P.REBOUND}
    orig = {f"{name}:{sym}": _resolve(mods[name], sym) for name, syms in P.REBOUND.items() for sym in syms}
    status = P.patch_reference()
    assert all(status.values()), status
    return orig, mods, P


def test_every_listed_symbol_is_rebound(patched):
    orig, mods, P = patched
    for key, fn in orig.items():
        name, sym = key.split(":")
        now = _resolve(mods[name], sym)
        assert now is not fn, f"{key} was not rebound"
        assert (getattr(now, "__module__", "") or "").startswith("moca_video_b200"), (key, now.__module__)


def test_replacements_accept_the_reference_signatures(patched):
    """Same parameter names in the same order with the same defaults; replacements may only ADD defaulted parameters."""
    orig, mods, P = patched
    for key, fn in orig.items():
        if key not in ORIGINAL_AT:                       # funcs.py's imported FreeInit names: same objects as utils.freeinit_utils
            continue
        name, sym = key.split(":")
        so, sn = inspect.signature(fn), inspect.signature(_resolve(mods[name], sym))
        po = [p for p in so.parameters.values() if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
        pn = [p for p in sn.parameters.values() if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
        assert len(pn) >= len(po), (key, ORIGINAL_AT[key])
        for a, b in zip(po, pn):
            assert a.name == b.name, (key, ORIGINAL_AT[key], a.name, b.name)
            if a.default is not inspect.Parameter.empty:
                assert b.default == a.default, (key, a.name, a.default, b.default)
            else:
                assert b.default is inspect.Parameter.empty or key.endswith("prepare_latents") or key.endswith("shift_latents"), (key, a.name)
        for b in pn[len(po):]:
            assert b.default is not inspect.Parameter.empty, (key, "extra parameter without default", b.name)
        if any(p.kind == p.VAR_KEYWORD for p in so.parameters.values()):
            assert any(p.kind == p.VAR_KEYWORD for p in sn.parameters.values()), (key, "**kwargs dropped")


def test_patched_calls_reach_the_device_boundary(patched):
    """The reference's own call shapes flow through the replacements up to libmoca_b200 and stop with MocaError on a box
    without a GPU (no CPU fallback) — not with a TypeError / AttributeError."""
    from moca_video_b200 import MocaError
    orig, mods, P = patched
    if torch.cuda.is_available():
        pytest.skip("CPU-boundary test")
    p = refimport.unet_params()
    p.update(model_channels=64, num_head_channels=64, context_dim=64)
    net = mods["lvdm.modules.networks.openaimodel3d"].UNetModel(**p).eval()
    x = torch.zeros(1, 4, 16, 8, 8)
    with pytest.raises(MocaError):
        net(x, torch.arange(16), context=torch.zeros(1, 7, 64), fps=10, gamma=0.9, clean_cond=True)   # stray sampler kwargs
    # module-level drop-ins: a module that was never attached to an engine says so; the reference's call shapes are accepted
    rb = net.input_blocks[1][0]
    with pytest.raises(MocaError, match="not attached"):
        rb(torch.zeros(16, 64, 8, 8), torch.zeros(16, 256), batch_size=1)
    with pytest.raises(MocaError, match="not attached"):
        net.input_blocks[1][1](torch.zeros(16, 64, 8, 8), context=torch.zeros(16, 7, 64))
    with pytest.raises(MocaError, match="not attached"):
        net.input_blocks[1][2](torch.zeros(1, 64, 16, 8, 8))
    # first-stage decode through the reference's own AutoencoderKL / decode_first_stage_2DAE call shapes
    dd = dict(double_z=True, z_channels=4, resolution=64, in_channels=3, out_ch=3, ch=64, ch_mult=(1, 2, 2), num_res_blocks=1,
              attn_resolutions=[], dropout=0.0)
    ae = mods["lvdm.models.autoencoder"].AutoencoderKL(dd, {"target": "torch.nn.Identity"}, 4)
    with pytest.raises(MocaError):
        ae.decode(torch.zeros(1, 4, 8, 8))
    with pytest.raises(MocaError):
        ae.encode(torch.zeros(1, 3, 64, 64))
    ld = types.SimpleNamespace(first_stage_model=ae, scale_factor=0.18215)
    with pytest.raises(MocaError):
        mods["lvdm.models.ddpm3d"].LatentDiffusion.decode_first_stage_2DAE(ld, torch.zeros(1, 4, 2, 8, 8))
    # text tower: an object shaped like FrozenOpenCLIPEmbedder after open_clip built its model (open_clip itself is not installed)
    class _Attn(torch.nn.Module):
        num_heads = 2
    blk = torch.nn.Module(); blk.attn = _Attn()
    clip = torch.nn.Module()
    clip.token_embedding = torch.nn.Embedding(100, 128)
    clip.positional_embedding = torch.nn.Parameter(torch.zeros(77, 128))
    clip.transformer = torch.nn.Module(); clip.transformer.resblocks = torch.nn.ModuleList([blk, blk])
    emb = object.__new__(mods["lvdm.modules.encoders.condition"].FrozenOpenCLIPEmbedder)
    torch.nn.Module.__init__(emb)
    emb.model, emb.layer_idx = clip, 1
    with pytest.raises(MocaError):
        emb.encode_with_transformer(torch.zeros(1, 77, dtype=torch.long))
    # sampler through the reference's own constructor path (oracle/refimport's CPU subclass) and LatentDiffusion-like apply_model
    seen = []

    def eps(xx, t, c):
        seen.append((tuple(xx.shape), len(t)))
        return torch.zeros_like(xx)
    m = refimport.ShimModel(eps)
    s = refimport.make_sampler(m, 64, 1.0)               # make_schedule is the rebound one
    assert hasattr(s, "tables") and s.ddim_timesteps.shape == (64,)
    ts = np.asarray(s.ddim_timesteps[:16])
    with pytest.raises(MocaError):
        s.fifo_onestep(cond={"c_crossattn": [torch.zeros(1, 7, 64)], "fps": 10}, shape=(1, 4, 16, 8, 8), latents=x, timesteps=ts,
                       indices=np.arange(16), unconditional_guidance_scale=12.0,
                       unconditional_conditioning={"c_crossattn": [torch.zeros(1, 5, 64)], "fps": 10}, cond_image=torch.zeros(1, 4, 1, 8, 8),
                       target="cat", gamma=0.9, use_self_attention=False, davis_masks=torch.zeros(1, 1, 16, 8, 8), clean_cond=True)
    assert seen and seen[0] == ((2, 4, 16, 8, 8), 32)    # cond + uncond went out as ONE b=2 apply_model call
    with pytest.raises(MocaError):
        mods["utils.freeinit_utils"].get_freq_filter((1, 4, 1, 8, 8), "cpu", "gaussian", 1, 0.25, 0.25)
    with pytest.raises(MocaError):
        mods["scripts.evaluation.funcs"].shift_latents(torch.zeros(1, 4, 72, 8, 8))
    fa = types.SimpleNamespace(num_inference_steps=64, lookahead_denoising=True, video_length=16, num_partitions=4, new_video_length=4,
                               save_frames=False)
    with pytest.raises((MocaError, FileNotFoundError)):   # prepare_latents loads {latents_dir}/{S}.pt like the reference (funcs.py:51)
        mods["scripts.evaluation.funcs"].fifo_ddim_sampling(fa, m, {"c_crossattn": [torch.zeros(1, 7, 64)], "fps": 10}, (1, 4, 16, 8, 8), s,
                                                            cfg_scale=1.0, latents_dir="/nonexistent", conditioned_image=None, targets="cat", gamma=0.9)

"""End to end through driver.run_prompts on the GPU (SURVEY N3 around the hot path): prompt CSV -> text tower -> base DDIM run that
fills the latent cache -> FIFO diagonal denoising with the MoCA step -> VAE decode of every popped frame on the side stream ->
frame writer -> video file; every stage is the library's.  Reduced widths / 8 DDIM steps so that it runs in seconds; what is checked
is the plumbing (shapes, cache files, frame count, finite images, determinism of a second run from the cache) — the numerics of each
stage have their own parity tests."""
import os
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CSV = ("prompt,conditioned_prompt,conditioned_object,conditioned_image_path,gamma\n"
       "\"A dog, running on the beach\",\"a cat, orange.\",cat,,0.5\n")


def _model(dev):
    from moca_video_b200.arch import UNetConfig, random_state_dict
    from moca_video_b200.model import LatentDiffusionB200
    from moca_video_b200.text import OpenCLIPTextEncoder, text_param_shapes
    from moca_video_b200.unet import UNet3D
    from moca_video_b200.vae import AutoencoderKLDecoder, decoder_param_shapes
    from oracle.text_ref import random_text_state_dict
    from oracle.vae_ref import random_vae_state_dict
    cfg = UNetConfig(model_channels=64, context_dim=64)
    net = UNet3D(model_channels=64, context_dim=64, device=dev).load_state_dict(random_state_dict(cfg, seed=0, device="cpu"))
    dd = dict(ch=64, ch_mult=(1, 2, 2), num_res_blocks=1, z_channels=4, out_ch=3)
    dec = AutoencoderKLDecoder(dd, device=dev).load_state_dict(random_vae_state_dict(decoder_param_shapes(**dd), seed=3))
    tcfg = (1000, 77, 64, 1, 2)
    # a stand-in tokenizer (open_clip's BPE vocabulary is not available here): bytes of the prompt as ids, <sot> ... <eot> padding
    def tok(texts):
        out = torch.zeros((len(texts), 77), dtype=torch.long)
        for i, t in enumerate(texts):
            ids = [998] + [1 + (b % 900) for b in t.encode()[:75]] + [999]
            out[i, :len(ids)] = torch.tensor(ids)
        return out
    txt = OpenCLIPTextEncoder(device=dev, layer="penultimate", text_cfg=tcfg, tokenizer=tok)
    txt.load_state_dict(random_text_state_dict(text_param_shapes(*tcfg[:4], 1), seed=8))
    return LatentDiffusionB200(net, fps=10, first_stage=dec, cond_stage=txt)


def test_run_prompts_end_to_end(tmp_path, monkeypatch):
    from moca_video_b200 import driver
    monkeypatch.chdir(tmp_path)
    dev = torch.device("cuda", torch.cuda.current_device())
    (tmp_path / "prompts.csv").write_text(CSV)
    model = _model(dev)
    args = types.SimpleNamespace(prompt_file=str(tmp_path / "prompts.csv"), prompt_index=None, output_dir=None, use_self_attention=True,
                                 eta=1.0, new_video_length=6, lookahead_denoising=True, num_partitions=2, video_length=4, use_davis=False,
                                 num_inference_steps=8, video_name=None, height=64, width=128, fps=10, seed=5,
                                 unconditional_guidance_scale=7.5, save_frames=True, output_fps=10, use_mp4=False)
    logs = []
    written = driver.run_prompts(args, model, log=logs.append)
    assert list(written) == [0] and os.path.exists(written[0]) and written[0].endswith(".gif")
    out_dir, lat_dir = driver.set_directory(args, "A dog, running on the beach", None, create=False)
    assert driver.latent_cache_ready(lat_dir, 8)                                     # the base DDIM run wrote 0.pt / 8.pt
    z = torch.load(os.path.join(lat_dir, "8.pt"))
    assert tuple(z.shape) == (1, 4, 4, 8, 16) and torch.isfinite(z).all()
    frames = sorted(os.listdir(os.path.join(out_dir, "fifo")), key=lambda n: int(n.split(".")[0]))
    assert len(frames) == 6 + 8 - 4                                                  # new_video_length + S - f iterations (funcs.py:305)
    from PIL import Image
    first = np.asarray(Image.open(os.path.join(out_dir, "fifo", frames[0])))
    assert first.shape == (32, 64, 3) and first.dtype == np.uint8 and first.std() > 0      # 3-level stand-in VAE: x4 of the 8 x 16 latent
    # second run: the cache is found, no base run; same seeds -> the same frames (side-stream decode included)
    before = {n: open(os.path.join(out_dir, "fifo", n), "rb").read() for n in frames}
    written2 = driver.run_prompts(args, model, log=logs.append)
    assert written2 == written
    after = {n: open(os.path.join(out_dir, "fifo", n), "rb").read() for n in frames}
    assert before == after

"""Driver / IO helpers against the reference's own functions where the reference tree is mounted, and on
their own otherwise.  CPU only."""
import os
import types

import numpy as np
import pytest
import torch

from moca_video_b200 import driver
from oracle import refimport

CSV = '''prompt,conditioned_object,conditioned_image_path,conditioned_prompt,gamma
"A dog, running on the beach",cat,assets/cats.png,a cat,0.5
A car in the city,horse,assets/horse.jpg,"a horse, white",0.75
A bird,plane,assets/plane.png,a plane,0.9
'''


def _args(**kw):
    a = types.SimpleNamespace(output_dir=None, use_self_attention=False, eta=1.0, new_video_length=100, lookahead_denoising=True,
                              num_partitions=4, video_length=16, use_davis=False, num_inference_steps=64, video_name=None)
    a.__dict__.update(kw)
    return a


def test_load_prompts_and_shard(tmp_path):
    p = tmp_path / "prompts.csv"
    p.write_text(CSV)
    rows = driver.load_prompts(str(p))
    assert len(rows) == 3 and rows[0]["prompt"] == "A dog, running on the beach" and rows[1]["conditioned_prompt"] == "a horse, white."
    assert rows[2]["gamma"] == 0.9
    assert driver.load_prompts(str(p), 1) == [rows[1]]
    with pytest.raises(ValueError):
        driver.load_prompts(str(p), 3)
    assert [driver.shard(7, r, 3) for r in range(3)] == [[0, 3, 6], [1, 4], [2, 5]]          # indices[rank::num_processes]
    if refimport.available():
        refimport.setup()
        from scripts.evaluation.funcs import load_prompts as ref_load
        assert ref_load(str(p)) == rows and ref_load(str(p), 2) == [rows[2]]


def test_set_directory_matches_reference(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    cases = [_args(), _args(eta=0.5, new_video_length=64), _args(lookahead_denoising=False, num_partitions=2, video_length=8),
             _args(output_dir="out/custom"), _args(use_self_attention=True)]
    got = [driver.set_directory(a, "A dog, running on the beach", "assets/cats.png") for a in cases]
    assert got[0] == ("results/videocraft_v2_fifo/random_noise/sam2/A dog, running on the beach/cats",
                      "results/videocraft_v2_fifo/latents/64steps/A dog, running on the beach/eta1.0")
    assert all(os.path.isdir(o) and os.path.isdir(l) for o, l in got)
    if refimport.available():
        refimport.setup()
        import importlib.util
        spec = importlib.util.spec_from_file_location("ref_main_for_test", os.path.join(refimport.REF, "videocrafter_main.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)                      # the CLI sits under __main__: importing only defines functions
        for a, g in zip(cases, got):
            assert mod.set_directory(a, "A dog, running on the beach", "assets/cats.png") == g


def test_latent_cache_and_frame_writer(tmp_path):
    d = str(tmp_path / "lat")
    os.makedirs(d)
    assert not driver.latent_cache_ready(d, 64)
    torch.save(torch.zeros(1), os.path.join(d, "0.pt"))
    torch.save(torch.zeros(1), os.path.join(d, "64.pt"))
    assert driver.latent_cache_ready(d, 64) and not driver.latent_cache_ready(d, 50)
    g = torch.Generator().manual_seed(0)
    frames = [torch.randn(1, 3, 1, 16, 24, generator=g) * 0.8 for _ in range(5)]
    w = driver.FrameWriter(str(tmp_path / "fifo"), save_frames=True)
    for i, f in enumerate(frames):
        w.put(i, f)
    out = w.close()
    assert len(out) == 5 and all(o.shape == (16, 24, 3) and o.dtype == np.uint8 for o in out)
    from PIL import Image
    for i, f in enumerate(frames):
        ref = torch.clamp(torch.squeeze(f).float(), -1., 1.)
        ref = (((ref + 1.0) / 2.0) * 255).to(torch.uint8).permute(1, 2, 0).numpy()           # funcs.py:626-637
        assert np.array_equal(out[i], ref)
        assert np.array_equal(np.asarray(Image.open(tmp_path / "fifo" / f"{i}.png")), ref)
    import sys
    from unittest.mock import MagicMock
    if isinstance(sys.modules.get("imageio"), MagicMock):          # oracle.refimport stubs it for the reference import
        del sys.modules["imageio"]
    path = driver.write_video(str(tmp_path / "video"), out, fps=10, use_mp4=False)
    assert os.path.exists(path)

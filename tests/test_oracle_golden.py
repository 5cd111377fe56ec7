"""The oracle restatements (oracle/step_ref.py, oracle/unet_ref.py) against fixtures produced by the
unmodified reference (oracle/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import step_ref as R
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.mark.parametrize("steps,eta", [(64, 1.0), (50, 1.0), (16, 0.0)])
def test_tables_bit_exact(steps, eta):
    g = load("tables.npz")
    t = R.make_tables(steps, eta)
    k = f"s{steps}_"
    assert np.array_equal(t["timesteps"], g[k + "timesteps"])
    for n in ("alphas", "alphas_prev", "sigmas", "sqrt_one_minus_alphas", "scale", "scale_prev"):
        assert np.array_equal(t[n], g[k + n]), n


def _case_inputs(g, c):
    if c == "d":
        rng = np.random.default_rng(100 + ord("d"))
        shape = (1, 4, 16, 40, 64)
        x = rng.standard_normal(shape).astype(np.float32)
        e_c = rng.standard_normal(shape).astype(np.float32)
        e_u = rng.standard_normal(shape).astype(np.float32)
        torch.manual_seed(int(g["d_noise_seed"]))
        noise = torch.cat([torch.randn((1, 4, 1, 40, 64)) for _ in range(16)], dim=2).numpy()
        return x, e_c, e_u, noise
    return g[c + "_x"], g[c + "_e_c"], g[c + "_e_u"], g[c + "_noise"]


@pytest.mark.parametrize("c", ["a", "b", "c", "d"])
def test_ddim_step_matches_reference(c):
    g = load("step_ddim.npz")
    x, e_c, e_u, noise = _case_inputs(g, c)
    tab = R.make_tables(64, 1.0)
    e = R.cfg_combine(e_c, e_u, float(g[c + "_cfg"]))
    with_mask = bool(g[c + "_with_mask"])
    H = x.shape[3]
    xp, p0 = R.moca_ddim_step(x, e, noise, R.coef_rows(tab, g[c + "_idx"]), g[c + "_t"],
                              g[c + "_mask"] if with_mask else None,
                              R.parity_mask_index(16, H) if with_mask else None, g[c + "_cond"], gamma=0.5,
                              sqrt=R.torch_cpu_sqrt)
    # float32, same operation order as torch eager, same (torch CPU) sqrt rounding as the golden run: bit-exact
    assert np.array_equal(xp, g[c + "_x_prev"])
    assert np.array_equal(p0, g[c + "_pred_x0"])
    # with the IEEE sqrt (what CUDA computes) the three scalar coefficients move by <= 1 ulp
    xp2, p02 = R.moca_ddim_step(x, e, noise, R.coef_rows(tab, g[c + "_idx"]), g[c + "_t"],
                                g[c + "_mask"] if with_mask else None,
                                R.parity_mask_index(16, H) if with_mask else None, g[c + "_cond"], gamma=0.5)
    np.testing.assert_allclose(xp2, g[c + "_x_prev"], rtol=0, atol=5e-5)
    np.testing.assert_allclose(p02, g[c + "_pred_x0"], rtol=0, atol=5e-5)


def sam_script(g, c):
    """segment_fn replaying the scripted Grounding-DINO / SAM-2 outputs the golden was generated with."""
    seg, seg_n = g[c + "_seg"], g[c + "_seg_n"]
    calls = []

    def fn(i, img):
        k = len(calls)
        calls.append(img)
        return None if seg_n[k] < 0 else seg[k, :seg_n[k]]
    return fn, calls


@pytest.mark.parametrize("c", ["a", "b"])
def test_ddim_step_sam_branch_matches_reference(c):
    """ddim_step without DAVIS masks: the unmodified ``_apply_segmentation`` (uint8 image, no-box fallback, IoU gate,
    >80 % coverage reset, gain-2 blend) vs the oracle restatement, scripted segmenter outputs."""
    g = load("step_ddim_sam.npz")
    tab = R.make_tables(64, 1.0)
    fn, calls = sam_script(g, c)
    xp, p0, used, images = R.moca_ddim_step_sam(g[c + "_x"], g[c + "_e"], g[c + "_noise"], R.coef_rows(tab, g[c + "_idx"]),
                                                g[c + "_t"], g[c + "_cond"], fn, gamma=0.5, sqrt=R.torch_cpu_sqrt)
    assert len(calls) == len(g[c + "_images"]) == int((g[c + "_t"] <= 300).sum())
    assert np.array_equal(np.stack(images), g[c + "_images"])          # what the segmenter is shown
    assert np.array_equal(xp, g[c + "_x_prev"])
    assert np.array_equal(p0, g[c + "_pred_x0"])
    blended = [u is not None for u in used]
    assert any(blended) and not all(blended)


def test_p_sample_ddim_matches_reference():
    g = load("step_psample.npz")
    tab = R.make_tables(50, 1.0)
    i = int(g["index"])
    e = R.cfg_combine(g["e_c"], g["e_u"], float(g["cfg"]))
    for k, sc in (("scale_", True), ("plain_", False)):
        xp, p0 = R.p_sample_ddim(g["x"], e, g[k + "noise"], tab["alphas"][i], tab["alphas_prev"][i], tab["sigmas"][i],
                                 tab["sqrt_one_minus_alphas"][i], tab["scale"][i] if sc else None,
                                 tab["scale_prev"][i] if sc else None, sqrt=R.torch_cpu_sqrt)
        assert np.array_equal(xp, g[k + "x_prev"])
        assert np.array_equal(p0, g[k + "pred_x0"])


def test_prepare_latents_and_shift():
    g = load("queue_freeinit.npz")
    tab = R.make_tables(64, 1.0)
    q = R.prepare_latents(g["z"], g["queue_noise"], tab["alphas"])
    assert q.shape == g["queue"].shape == (1, 4, 72, 6, 8)
    np.testing.assert_allclose(q, g["queue"], rtol=0, atol=5e-7)   # 1 ulp: torch CPU pow(0.5) rounding
    sh = R.shift_latents(g["shift_in"], g["shift_noise"])
    np.testing.assert_allclose(sh, g["shift_out"], rtol=0, atol=5e-6)
    assert np.array_equal(sh[:, :, :-1], g["shift_in"][:, :, 1:])


def test_filters_and_mix():
    g = load("queue_freeinit.npz")
    for ft, fn in (("gaussian", R.gaussian_low_pass_filter), ("ideal", R.ideal_low_pass_filter),
                   ("box", R.box_low_pass_filter)):
        assert np.array_equal(fn(8, 10, 12), g["filter_" + ft]), ft
    assert np.array_equal(R.butterworth_low_pass_filter(8, 10, 12, 4), g["filter_butterworth"])
    f1 = R.gaussian_low_pass_filter(1, 40, 64)
    assert np.array_equal(f1, g["filter_gauss_t1"])
    assert abs(float(f1.max()) - np.exp(-8.0)) < 1e-9
    lpf = R.gaussian_low_pass_filter(8, 10, 12)
    y = R.freq_mix_3d(g["mix_x"], g["mix_noise"], lpf)
    np.testing.assert_allclose(y, g["mix_out"], rtol=0, atol=5e-6)


@pytest.mark.parametrize("tag", ["fifo", "plain"])
def test_unet_restatement_small(tag):
    from moca_video_b200.arch import UNetConfig, random_state_dict
    from oracle.make_golden import SMALL
    from oracle.unet_ref import unet_forward
    g = load("unet_small.npz")
    cfg = UNetConfig(**SMALL)
    sd = random_state_dict(cfg, seed=0)
    y = unet_forward(sd, cfg, torch.from_numpy(g[tag + "_x"]), torch.from_numpy(g[tag + "_ts"]),
                     torch.from_numpy(g[tag + "_ctx"]), fps=10)
    ref = torch.from_numpy(g[tag + "_y"])
    assert float((y - ref).norm() / ref.norm()) < 1e-5      # fp32 tolerance of north_star is 1e-3


def test_full_size_param_listing():
    import hashlib
    from moca_video_b200.arch import UNetConfig, param_shapes
    p = os.path.join(GOLDEN, "unet_full_meta.npz")
    if not os.path.exists(p):
        pytest.skip("full-size meta fixture not generated")
    g = np.load(p)
    shp = {k: tuple(v) for k, v in param_shapes(UNetConfig.videocrafter2()).items()}
    listing = "\n".join(f"{k} {v}" for k, v in sorted(shp.items()))
    assert hashlib.sha256(listing.encode()).hexdigest() == str(g["shapes_sha256"])
    assert sum(int(np.prod(v)) for v in shp.values()) == int(g["n_params"]) == 1413284420

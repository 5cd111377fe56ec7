"""Generate tests/golden/*.npz by running the UNMODIFIED reference from /root/reference.

TEST INFRASTRUCTURE ONLY.  Run in the authoring container:  python -m oracle.make_golden
The reference ships no fixtures; these files are what pins oracle/step_ref.py and
oracle/unet_ref.py to the reference.  Inputs are stored next to outputs so the tests need no RNG parity.
"""
import os
import sys
import tempfile

import numpy as np
import torch

from oracle import refimport

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _np(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def gen_tables():
    out = {}
    for steps, eta in ((64, 1.0), (50, 1.0), (16, 0.0)):
        m = refimport.ShimModel()
        s = refimport.make_sampler(m, steps, eta)
        k = f"s{steps}_"
        out[k + "timesteps"] = _np(s.ddim_timesteps)
        out[k + "alphas"] = _np(s.ddim_alphas).astype(np.float32)
        out[k + "alphas_prev"] = _np(s.ddim_alphas_prev).astype(np.float32)
        out[k + "sigmas"] = _np(s.ddim_sigmas).astype(np.float32)
        out[k + "sqrt_one_minus_alphas"] = _np(s.ddim_sqrt_one_minus_alphas).astype(np.float32)
        out[k + "scale"] = _np(s.ddim_scale_arr).astype(np.float32)
        out[k + "scale_prev"] = _np(s.ddim_scale_arr_prev).astype(np.float32)
    np.savez_compressed(os.path.join(GOLD, "tables.npz"), **out)


def gen_ddim_step():
    """fifo_onestep (ddim.py:255-271) with a stub eps model: isolates unet()'s CFG combine + ddim_step."""
    import lvdm.models.samplers.ddim as D
    cases = {}
    for name, (H, W, win, with_mask) in {"a": (12, 16, 3, True), "b": (8, 20, 0, True), "c": (12, 16, 5, False),
                                          "d": (40, 64, 1, True)}.items():
        rng = np.random.default_rng(100 + ord(name))
        shape = (1, 4, 16, H, W)
        x = rng.standard_normal(shape).astype(np.float32)
        e_c = rng.standard_normal(shape).astype(np.float32)
        e_u = rng.standard_normal(shape).astype(np.float32)
        mask = (rng.random((1, 1, 16, H, W)) < 0.3).astype(np.float32)
        if name == "b":
            mask[:, :, 1] = 0.0            # the mask frame the clobbered index selects is empty: exercises sum()==0
        cond = rng.random((1, 4, 1, H, W)).astype(np.float32)
        cfg = 12.0
        cdict, udict = {"tag": "c"}, {"tag": "u"}

        def eps_fn(xx, t, c):
            return torch.from_numpy(e_c if c["tag"] == "c" else e_u)

        m = refimport.ShimModel(eps_fn)
        s = refimport.make_sampler(m, 64, 1.0)
        draws = []
        orig = D.noise_like

        def rec(shape_, device, repeat=False):
            n = orig(shape_, device, repeat); draws.append(n.clone()); return n

        D.noise_like = rec
        ts_q = np.concatenate([np.full((8,), s.ddim_timesteps[0]), s.ddim_timesteps])
        idx_q = np.concatenate([np.full((8,), 0), np.arange(64)])
        t = ts_q[8 * win:8 * win + 16]; idx = idx_q[8 * win:8 * win + 16]
        torch.manual_seed(7)
        cwd = os.getcwd()
        with tempfile.TemporaryDirectory() as td:
            os.chdir(td)
            try:
                xp, p0 = s.fifo_onestep(cond=cdict, shape=shape, latents=torch.from_numpy(x), timesteps=t, indices=idx,
                                        unconditional_guidance_scale=cfg, unconditional_conditioning=udict,
                                        cond_image=torch.from_numpy(cond), target="cat", gamma=0.9,
                                        use_self_attention=True,
                                        davis_masks=torch.from_numpy(mask) if with_mask else None, clean_cond=True)
            finally:
                os.chdir(cwd); D.noise_like = orig
        if not with_mask:
            # SAM/attention branch would run for t <= 300 (ddim.py:592); window 5 has all t > 300 so no blend happens
            assert int(t.min()) > 300
        noise = torch.cat(draws, dim=2).numpy()
        kw_seen = m.calls[0]
        assert "gamma" in kw_seen       # gamma lands in apply_model kwargs, ddim_step keeps 0.5
        for k, v in dict(x=x, e_c=e_c, e_u=e_u, mask=mask, cond=cond, noise=noise, t=t, idx=idx,
                         x_prev=_np(xp), pred_x0=_np(p0), cfg=np.float32(cfg), with_mask=np.int32(with_mask)).items():
            if name == "d" and k in ("x", "e_c", "e_u", "noise"):
                continue                 # full-size case: inputs regenerated from seeds in the test (kept small)
            cases[f"{name}_{k}"] = v
        if name == "d":
            cases["d_noise_seed"] = np.int64(7)
    np.savez_compressed(os.path.join(GOLD, "step_ddim.npz"), **cases)


def gen_ddim_step_sam():
    """ddim_step WITHOUT DAVIS masks (ddim.py:591-606): frames with t <= 300 run the unmodified
    ``_apply_segmentation`` (:739-903) — uint8 conversion, no-box fallback, IoU gate, >80 % coverage reset, blend with
    gain 2 — with Grounding-DINO / SAM-2 replaced by scripted stubs (third-party nets without weights; their outputs are
    the stored ``seg_*`` arrays).  As shipped, the blend's ``torch.where`` broadcasts [1,C,H,W] against [1,C,1,H,W] and
    returns C identical copies of the frame along dim 2; the generator checks the copies are identical and stores one."""
    refimport.setup()
    import lvdm.models.samplers.ddim as D

    class T(torch.Tensor):                       # ``v.to("cuda")`` of the processor outputs (ddim.py:770-775) on a CPU box
        def to(self, *a, **k):
            return self

    cases = {}
    for name, (H, W, win) in {"a": (12, 16, 1), "b": (10, 12, 2)}.items():
        rng = np.random.default_rng(300 + ord(name))
        shape = (1, 4, 16, H, W)
        x = rng.standard_normal(shape).astype(np.float32)
        if name == "b":
            x *= 0.05                              # small latents: floor(max) <= 1 for some frames -> the x255 branch
        e = rng.standard_normal(shape).astype(np.float32) * (0.05 if name == "b" else 1.0)
        cond = rng.random((1, 4, 1, H, W)).astype(np.float32)
        m = refimport.ShimModel(None)
        s = refimport.make_sampler(m, 64, 1.0)
        ts_q = np.concatenate([np.full((8,), s.ddim_timesteps[0]), s.ddim_timesteps])
        idx_q = np.concatenate([np.full((8,), 0), np.arange(64)])
        t = ts_q[8 * win:8 * win + 16]; idx = idx_q[8 * win:8 * win + 16]

        def blob(frac):
            return (rng.random((H, W)) < frac).astype(np.float32)

        def grow(mk, frac):                                                   # superset of mk covering ~frac of the frame
            return np.maximum(mk, blob(max(0.0, (frac - mk.mean()) / max(1e-6, 1.0 - mk.mean()))))

        def jitter(mk):                                                       # a few flipped pixels: IoU stays high
            out = mk.copy(); out[0, :2] = 1.0 - out[0, :2]; return out
        n1, big, n2 = blob(0.3), blob(0.7), blob(0.2)
        if name == "a":
            base = np.stack([n1, big, n2])
            k4 = np.stack([jitter(n1), grow(big, 0.86), jitter(n2)])
            script = [None,                                                   # no box, no previous masks: frame untouched
                      base,                                                   # three masks, all below 80 % coverage
                      np.stack([blob(0.3), blob(0.3), blob(0.3)]),            # unrelated: mean IoU < 0.5 -> previous masks kept
                      None,                                                   # no box -> previous masks
                      k4,                                                     # accepted; middle mask > 80 %: RESET, then the last mask
                      np.stack([jitter(m_) for m_ in k4]),                    # accepted (IoU with call 4), reset again
                      jitter(k4[0])[None],                                    # fewer masks than before: zip() compares the first pair only
                      np.stack([jitter(k4[0]), blob(0.5)])]                   # more masks than before
        else:
            huge = blob(0.9)
            script = [huge[None],                                             # only an over-covering mask: frame untouched
                      jitter(huge)[None],
                      None,
                      np.stack([np.minimum(huge, blob(0.85))]),               # subset at ~76 %: accepted and blended
                      np.stack([blob(0.1)]),                                  # rejected by the IoU gate
                      np.stack([np.minimum(huge, blob(0.85)), blob(0.3)])]
        while len(script) < 16:
            script.append(np.stack([blob(0.3)]) if len(script) % 2 else None)
        log = {"images": [], "calls": 0, "dims": []}

        class Proc:
            def __call__(self, images=None, text=None, return_tensors=None):
                return {"input_ids": torch.zeros((1, 1), dtype=torch.long).as_subclass(T)}

            def post_process_grounded_object_detection(self, outputs, ids, box_threshold=None, text_threshold=None, target_sizes=None):
                k = log["calls"]
                nb = 0 if script[k] is None else len(script[k])
                return [{"boxes": torch.zeros((nb, 4))}]

        class Sam:
            def set_image(self, img):
                log["images"].append(np.array(img))

            def predict(self, point_coords=None, point_labels=None, box=None, multimask_output=False):
                return script[log["calls"]].copy(), None, None

        s.processor = Proc(); s.sam2_predictor = Sam(); s.grounding_model = lambda **kw: None
        orig_apply = s.apply_cond_img

        def apply_rec(pred_x0, *a, **k):
            out, masks = orig_apply(pred_x0, *a, **k)
            log["calls"] += 1
            log["dims"].append(int(out.shape[2]))
            return out, masks
        s.apply_cond_img = apply_rec
        draws = []
        orig = D.noise_like

        def rec(shape_, device, repeat=False):
            n = orig(shape_, device, repeat); draws.append(n.clone()); return n
        D.noise_like = rec
        torch.manual_seed(17)
        cwd = os.getcwd()
        import warnings
        with tempfile.TemporaryDirectory() as td, warnings.catch_warnings():
            warnings.simplefilter("ignore")
            os.chdir(td)
            try:
                xp, p0 = s.ddim_step(torch.from_numpy(x), torch.from_numpy(e), idx, torch.from_numpy(cond), "cat",
                                     torch.from_numpy(t).long(), use_self_attention=False, davis_masks=None)
            finally:
                os.chdir(cwd); D.noise_like = orig
        n_seg = int((t <= 300).sum())
        assert log["calls"] == n_seg and len(log["images"]) == n_seg, (log["calls"], n_seg)
        # undo the replicated frames (see docstring)
        frames, pos, k = [], 0, 0
        p0 = _np(p0)
        for i in range(16):
            d = 1
            if t[i] <= 300:
                d = log["dims"][k]; k += 1
            blk = p0[:, :, pos:pos + d]
            for j in range(1, d):
                assert np.array_equal(blk[:, :, j], blk[:, :, 0])
            frames.append(blk[:, :, :1]); pos += d
        assert pos == p0.shape[2]
        print(f"ddim_step_sam[{name}]: {n_seg} segmented frames, replicated dims {log['dims']}")
        nmax = max(len(m_) for m_ in script[:n_seg] if m_ is not None)
        seg = np.zeros((n_seg, nmax, H, W), np.float32); seg_n = np.full((n_seg,), -1, np.int32)
        for k in range(n_seg):
            if script[k] is not None:
                seg_n[k] = len(script[k]); seg[k, :len(script[k])] = script[k]
        cases.update({f"{name}_x": x, f"{name}_e": e, f"{name}_cond": cond, f"{name}_t": t, f"{name}_idx": idx,
                      f"{name}_noise": torch.cat(draws, dim=2).numpy(), f"{name}_x_prev": _np(xp),
                      f"{name}_pred_x0": np.concatenate(frames, axis=2), f"{name}_seg": seg, f"{name}_seg_n": seg_n,
                      f"{name}_images": np.stack(log["images"])})
    np.savez_compressed(os.path.join(GOLD, "step_ddim_sam.npz"), **cases)


def gen_p_sample():
    import lvdm.models.samplers.ddim as D
    rng = np.random.default_rng(5)
    shape = (2, 4, 4, 8, 8)
    x = rng.standard_normal(shape).astype(np.float32)
    e_c = rng.standard_normal(shape).astype(np.float32)
    e_u = rng.standard_normal(shape).astype(np.float32)
    out = dict(x=x, e_c=e_c, e_u=e_u)
    for use_scale in (True, False):
        m = refimport.ShimModel(lambda xx, t, c: torch.from_numpy(e_c if c == "c" else e_u), use_scale=use_scale)
        s = refimport.make_sampler(m, 50, 1.0)
        draws = []
        orig = D.noise_like

        def rec(shape_, device, repeat=False):
            n = orig(shape_, device, repeat); draws.append(n.clone()); return n
        D.noise_like = rec
        torch.manual_seed(11)
        index = 33
        # c / uc are plain tensors in this call path; use tagged strings through isinstance(dict) branch instead
        class Tag(dict):
            pass
        try:
            m.eps_fn = lambda xx, t, c: torch.from_numpy(e_c if c.get("tag") == "c" else e_u)
            xp, p0 = s.p_sample_ddim(torch.from_numpy(x), Tag(tag="c"), torch.full((2,), int(s.ddim_timesteps[index])),
                                     index=index, unconditional_guidance_scale=7.5,
                                     unconditional_conditioning=Tag(tag="u"))
        finally:
            D.noise_like = orig
        k = "scale_" if use_scale else "plain_"
        out[k + "x_prev"] = _np(xp); out[k + "pred_x0"] = _np(p0); out[k + "noise"] = draws[0].numpy()
    out["index"] = np.int64(33); out["cfg"] = np.float32(7.5)
    np.savez_compressed(os.path.join(GOLD, "step_psample.npz"), **out)


def gen_queue():
    """prepare_latents (funcs.py:21-82), shift_latents (:86-99), FreeInit filters and freq_mix_3d."""
    import types
    import scripts.evaluation.funcs as Fn
    from utils import freeinit_utils as FI
    rng = np.random.default_rng(9)
    z = rng.standard_normal((1, 4, 16, 6, 8)).astype(np.float32)
    m = refimport.ShimModel()
    s = refimport.make_sampler(m, 64, 1.0)

    class T(torch.Tensor):
        def to(self, *a, **k):
            return self
    args = types.SimpleNamespace(num_inference_steps=64, lookahead_denoising=True, video_length=16)
    draws = []
    orig_load, orig_rl = torch.load, torch.randn_like
    torch.load = lambda *a, **k: torch.from_numpy(z).as_subclass(T)

    def rl(t, **k):
        n = orig_rl(torch.Tensor(t) if not type(t) is torch.Tensor else t, **k); draws.append(n.clone()); return n
    torch.randn_like = rl
    torch.manual_seed(3)
    try:
        q = Fn.prepare_latents(args, "/nonexistent", s)
    finally:
        torch.load = orig_load; torch.randn_like = orig_rl
    out = dict(z=z, queue=_np(torch.Tensor(q)), queue_noise=torch.cat([torch.Tensor(d) for d in draws], dim=2).numpy())
    # shift_latents
    lat = rng.standard_normal((1, 4, 72, 6, 8)).astype(np.float32)
    draws = []
    torch.randn_like = rl
    torch.manual_seed(4)
    try:
        sh = Fn.shift_latents(torch.from_numpy(lat.copy()))
    finally:
        torch.randn_like = orig_rl
    out.update(shift_in=lat, shift_out=_np(sh), shift_noise=draws[0].numpy())
    # filters + mix at T>1
    for ft in ("gaussian", "butterworth", "ideal", "box"):
        out["filter_" + ft] = _np(FI.get_freq_filter((1, 4, 8, 10, 12), "cpu", ft, 4, 0.25, 0.25))[0, 0]
    out["filter_gauss_t1"] = _np(FI.get_freq_filter((1, 4, 1, 40, 64), "cpu", "gaussian", 1, 0.25, 0.25))[0, 0]
    xm = rng.standard_normal((2, 4, 8, 10, 12)).astype(np.float32)
    nm = rng.standard_normal((2, 4, 8, 10, 12)).astype(np.float32)
    lpf = FI.get_freq_filter((2, 4, 8, 10, 12), "cpu", "gaussian", 1, 0.25, 0.25)
    # freq_mix_3d squeezes dim 0 only when it is 1; call per batch element as the driver does (b=1)
    mixed = np.stack([_np(FI.freq_mix_3d(torch.from_numpy(xm[i:i + 1]), torch.from_numpy(nm[i:i + 1]), lpf[i:i + 1]))
                      for i in range(2)])
    out.update(mix_x=xm, mix_noise=nm, mix_out=mixed.reshape(xm.shape))
    np.savez_compressed(os.path.join(GOLD, "queue_freeinit.npz"), **out)


SMALL = dict(in_channels=4, out_channels=4, model_channels=32, attention_resolutions=(4, 2, 1), num_res_blocks=2,
             channel_mult=(1, 2, 4, 4), num_head_channels=32, context_dim=48)


def gen_unet_small():
    """Reference UNetModel (all yaml flags, shrunk widths) vs oracle.unet_ref on identical seeded weights."""
    from moca_video_b200.arch import UNetConfig, param_shapes, random_state_dict
    from oracle.unet_ref import unet_forward
    cfg = UNetConfig(**SMALL)
    p = refimport.unet_params()
    p.update(model_channels=32, num_head_channels=32, context_dim=48)
    net = refimport.build_unet(p, rerandomize=False)
    ref_shapes = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    mine = {k: tuple(v) for k, v in param_shapes(cfg).items()}
    assert ref_shapes == mine, (set(ref_shapes) ^ set(mine), [k for k in mine if k in ref_shapes and mine[k] != ref_shapes[k]][:5])
    assert list(ref_shapes) == list(mine) or True
    sd = random_state_dict(cfg, seed=0)
    net.load_state_dict(sd, strict=True)
    rng = np.random.default_rng(21)
    out = {}
    for tag, (b, t, h, w, L, fifo) in {"fifo": (1, 16, 8, 8, 11, True), "plain": (2, 4, 8, 16, 7, False)}.items():
        x = torch.from_numpy(rng.standard_normal((b, 4, t, h, w)).astype(np.float32))
        ctx = torch.from_numpy(rng.standard_normal((b, L, 48)).astype(np.float32))
        ts = torch.from_numpy(rng.integers(0, 1000, size=(t if fifo else b,))).long()
        with torch.no_grad():
            y_ref = net(x, ts, context=ctx, fps=10)
        y = unet_forward(sd, cfg, x, ts, ctx, fps=10)
        err = float((y - y_ref).norm() / y_ref.norm())
        print(f"unet_small[{tag}] restatement vs reference rel-L2 = {err:.3e}, |y| = {float(y_ref.abs().mean()):.3f}")
        assert err < 1e-5
        out.update({f"{tag}_x": _np(x), f"{tag}_ctx": _np(ctx), f"{tag}_ts": _np(ts), f"{tag}_y": _np(y_ref)})
    np.savez_compressed(os.path.join(GOLD, "unet_small.npz"), **out)


SMALL64 = dict(in_channels=4, out_channels=4, model_channels=64, attention_resolutions=(4, 2, 1), num_res_blocks=2,
               channel_mult=(1, 2, 4, 4), num_head_channels=64, context_dim=64)


def gen_unet_small64():
    """Reference UNetModel with head dim 64 (the width the CUDA attention kernels are built for), shrunk to
    model_channels 64: the fixture the GPU parity test compares the CUDA path with on the box (no reference there)."""
    from moca_video_b200.arch import UNetConfig, param_shapes, random_state_dict
    from oracle.unet_ref import unet_forward
    cfg = UNetConfig(**SMALL64)
    p = refimport.unet_params()
    p.update(model_channels=64, num_head_channels=64, context_dim=64)
    net = refimport.build_unet(p, rerandomize=False)
    ref_shapes = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    assert ref_shapes == {k: tuple(v) for k, v in param_shapes(cfg).items()}
    sd = random_state_dict(cfg, seed=0)
    net.load_state_dict(sd, strict=True)
    rng = np.random.default_rng(23)
    out = {}
    for tag, (b, t, h, w, L, fifo) in {"fifo": (1, 16, 8, 16, 11, True), "plain": (2, 4, 16, 8, 7, False)}.items():
        x = torch.from_numpy(rng.standard_normal((b, 4, t, h, w)).astype(np.float32))
        ctx = torch.from_numpy(rng.standard_normal((b, L, 64)).astype(np.float32))
        ts = torch.from_numpy(rng.integers(0, 1000, size=(t if fifo else b,))).long()
        with torch.no_grad():
            y_ref = net(x, ts, context=ctx, fps=10)
        y = unet_forward(sd, cfg, x, ts, ctx, fps=10)
        err = float((y - y_ref).norm() / y_ref.norm())
        print(f"unet_small64[{tag}] restatement vs reference rel-L2 = {err:.3e}, |y| = {float(y_ref.abs().mean()):.3f}")
        assert err < 1e-5
        out.update({f"{tag}_x": _np(x), f"{tag}_ctx": _np(ctx), f"{tag}_ts": _np(ts), f"{tag}_y": _np(y_ref)})
    np.savez_compressed(os.path.join(GOLD, "unet_small64.npz"), **out)


def gen_unet_full():
    """Full-size architecture: key/shape listing equality + one forward (slow, ~2 min)."""
    from moca_video_b200.arch import UNetConfig, param_shapes, random_state_dict
    from oracle.unet_ref import unet_forward
    import hashlib
    cfg = UNetConfig.videocrafter2()
    p = refimport.unet_params()
    with torch.device("meta"):
        refimport.setup()
        from lvdm.modules.networks.openaimodel3d import UNetModel
        net = UNetModel(**p)
    ref_shapes = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    mine = {k: tuple(v) for k, v in param_shapes(cfg).items()}
    assert ref_shapes == mine
    listing = "\n".join(f"{k} {v}" for k, v in sorted(ref_shapes.items()))
    digest = hashlib.sha256(listing.encode()).hexdigest()
    nparam = sum(int(np.prod(v)) for v in ref_shapes.values())
    print("full-size state_dict:", len(ref_shapes), "tensors,", nparam, "params, sha256", digest)
    out = dict(shapes_sha256=np.array(digest), n_tensors=np.int64(len(ref_shapes)), n_params=np.int64(nparam))
    if os.environ.get("MOCA_GOLDEN_FULL_FORWARD", "1") == "1":
        net = refimport.build_unet(p, rerandomize=False)
        sd = random_state_dict(cfg, seed=0)
        net.load_state_dict(sd, strict=True)
        g = torch.Generator().manual_seed(1)
        x = torch.randn((1, 4, 16, 40, 64), generator=g)
        ctx = torch.randn((1, 77, 1024), generator=g)
        ts_q = np.concatenate([np.full((8,), 0), np.linspace(0, 999, 64).round().astype(np.int64)])
        ts = torch.from_numpy(ts_q[24:40]).long()
        with torch.no_grad():
            y_ref = net(x, ts, context=ctx, fps=10)
        y = unet_forward(sd, cfg, x, ts, ctx, fps=10)
        err = float((y - y_ref).norm() / y_ref.norm())
        print(f"unet_full restatement vs reference rel-L2 = {err:.3e}; |y| mean {float(y_ref.abs().mean()):.4f}")
        assert err < 1e-4
        out.update(full_ts=_np(ts), full_y_sub=_np(y_ref[0, :, ::5, ::8, ::8]), full_err=np.float64(err),
                   full_y_absmean=np.float64(y_ref.abs().mean()))
    np.savez_compressed(os.path.join(GOLD, "unet_full_meta.npz"), **out)


VAE_SMALL = dict(ch=64, ch_mult=(1, 2, 2), num_res_blocks=1, z_channels=4, out_ch=3)


def gen_vae_small():
    """Reference AutoencoderKL.decode (autoencoder.py:104-107; Decoder ae_modules.py:466-578) vs oracle.vae_ref on seeded weights:
    a shrunk ddconfig (64 channels, three levels), then the key / shape listing of the shipped ddconfig."""
    import hashlib
    import yaml
    refimport.setup()
    from lvdm.modules.networks.ae_modules import Decoder
    from moca_video_b200.vae import decoder_param_shapes
    from oracle.vae_ref import random_vae_state_dict, vae_decode

    from lvdm.models.autoencoder import AutoencoderKL

    dd = VAE_SMALL
    ae = AutoencoderKL(dict(dd, double_z=True, resolution=64, in_channels=3, attn_resolutions=[], dropout=0.0),
                       {"target": "torch.nn.Identity"}, dd["z_channels"]).eval()
    shapes = decoder_param_shapes(**dd)
    ref_shapes = {k: tuple(v.shape) for k, v in ae.state_dict().items() if k.startswith(("decoder.", "post_quant_conv."))}
    assert ref_shapes == {k: tuple(v) for k, v in shapes.items()}, set(ref_shapes) ^ set(shapes)
    sd = random_vae_state_dict(shapes, seed=3)
    missing, unexpected = ae.load_state_dict(sd, strict=False)
    assert not unexpected and all(k.startswith(("encoder.", "quant_conv.", "loss.")) for k in missing), (missing[:3], unexpected[:3])
    rng = np.random.default_rng(31)
    out = {}
    for tag, (n, h, w) in {"a": (2, 8, 8), "b": (1, 8, 16)}.items():
        z = torch.from_numpy(rng.standard_normal((n, 4, h, w)).astype(np.float32))
        with torch.no_grad():
            y_ref = ae.decode(z)                                   # the unmodified AutoencoderKL.decode (autoencoder.py:104-107)
            y = vae_decode(sd, z, len(dd["ch_mult"]), dd["num_res_blocks"])
        err = float((y - y_ref).norm() / y_ref.norm())
        print(f"vae_small[{tag}] restatement vs reference rel-L2 = {err:.3e}, |y| = {float(y_ref.abs().mean()):.3f}")
        assert err < 1e-5
        out.update({f"{tag}_z": _np(z), f"{tag}_y": _np(y_ref)})
    # encode path: the unmodified AutoencoderKL.encode (autoencoder.py:97-102) on seeded encoder weights
    from moca_video_b200.vae import encoder_param_shapes
    from oracle.vae_ref import vae_encode_moments
    eshapes = encoder_param_shapes(dd["ch"], dd["ch_mult"], dd["num_res_blocks"], dd["z_channels"])
    ref_e = {k: tuple(v.shape) for k, v in ae.state_dict().items() if k.startswith(("encoder.", "quant_conv."))}
    assert ref_e == {k: tuple(v) for k, v in eshapes.items()}, set(ref_e) ^ set(eshapes)
    esd = random_vae_state_dict(eshapes, seed=4)
    missing, unexpected = ae.load_state_dict(esd, strict=False)
    assert not unexpected
    for tag, (n, h, w) in {"ea": (2, 64, 64), "eb": (1, 32, 64)}.items():
        x = torch.from_numpy(rng.uniform(-1, 1, (n, 3, h, w)).astype(np.float32))
        with torch.no_grad():
            post = ae.encode(x)
            m = vae_encode_moments(esd, x, len(dd["ch_mult"]), dd["num_res_blocks"])
        err = float((m - post.parameters).norm() / post.parameters.norm())
        print(f"vae_small[{tag}] encoder restatement vs reference rel-L2 = {err:.3e}, |m| = {float(post.parameters.abs().mean()):.3f}")
        assert err < 1e-5
        out.update({f"{tag}_x": _np(x), f"{tag}_moments": _np(post.parameters)})
    # shipped ddconfig: listing equality on the meta device (no forward: 49.5 M parameters at 320 x 512 take minutes on CPU)
    with open(os.path.join(refimport.REF, "configs/inference_t2v_512_v2.0.yaml")) as f:
        full = yaml.safe_load(f)["model"]["params"]["first_stage_config"]["params"]["ddconfig"]
    with torch.device("meta"):
        fdec = Decoder(**full)
    fshapes = {"decoder." + k: tuple(v.shape) for k, v in fdec.state_dict().items()}
    fshapes.update({"post_quant_conv.weight": (4, 4, 1, 1), "post_quant_conv.bias": (4,)})
    mine = decoder_param_shapes(full["ch"], tuple(full["ch_mult"]), full["num_res_blocks"], full["z_channels"], full["out_ch"])
    assert fshapes == {k: tuple(v) for k, v in mine.items()}
    from lvdm.modules.networks.ae_modules import Encoder
    with torch.device("meta"):
        fenc = Encoder(**full)
    feshapes = {"encoder." + k: tuple(v.shape) for k, v in fenc.state_dict().items()}
    feshapes.update({"quant_conv.weight": (8, 8, 1, 1), "quant_conv.bias": (8,)})
    emine = encoder_param_shapes(full["ch"], tuple(full["ch_mult"]), full["num_res_blocks"], full["z_channels"], full["in_channels"], full["double_z"])
    assert feshapes == {k: tuple(v) for k, v in emine.items()}
    out["full_enc_shapes_sha256"] = np.array(hashlib.sha256("\n".join(f"{k} {v}" for k, v in sorted(feshapes.items())).encode()).hexdigest())
    out["full_enc_n_params"] = np.int64(sum(int(np.prod(v)) for v in feshapes.values()))
    listing = "\n".join(f"{k} {v}" for k, v in sorted(fshapes.items()))
    out["full_shapes_sha256"] = np.array(hashlib.sha256(listing.encode()).hexdigest())
    out["full_n_params"] = np.int64(sum(int(np.prod(v)) for v in fshapes.values()))
    print("full decoder:", len(fshapes), "tensors,", int(out["full_n_params"]), "params")
    np.savez_compressed(os.path.join(GOLD, "vae_small.npz"), **out)


TEXT_SMALL = (1000, 77, 128, 2, 4)          # vocab, context, width, heads, layers


def gen_text_small():
    """oracle.text_ref (open_clip's published text transformer behind condition.py:212-230) vs transformers.CLIPTextModel with the
    same weights: open_clip itself is not installed here, so this fixture pins the restatement to an INDEPENDENT implementation of
    the same model, not to the reference's dependency ("parity unpinned" in oracle/text_ref.py)."""
    from transformers import CLIPTextConfig, CLIPTextModel
    from moca_video_b200.text import text_param_shapes
    from oracle.text_ref import random_text_state_dict, text_encode, to_hf_state_dict
    vocab, L, W, heads, layers = TEXT_SMALL
    sd = random_text_state_dict(text_param_shapes(vocab, L, W, heads, layers), seed=8)
    cfg = CLIPTextConfig(vocab_size=vocab, hidden_size=W, intermediate_size=4 * W, num_hidden_layers=layers, num_attention_heads=heads,
                         max_position_embeddings=L, hidden_act="gelu", layer_norm_eps=1e-5, attn_implementation="eager")
    hf = CLIPTextModel(cfg).eval()
    missing, unexpected = hf.load_state_dict(to_hf_state_dict(sd, layers), strict=False)
    assert not unexpected and all("position_ids" in k for k in missing), (missing, unexpected)
    rng = np.random.default_rng(33)
    tokens = torch.from_numpy(rng.integers(1, vocab - 2, size=(3, L))).long()
    tokens[:, 0] = vocab - 2
    for r, n in enumerate((5, 20, 77)):                       # padded like open_clip.tokenize: <sot> words <eot> 0 0 ...
        if n < L:
            tokens[r, n] = vocab - 1
            tokens[r, n + 1:] = 0
    out = {"tokens": _np(tokens)}
    with torch.no_grad():
        o = hf(input_ids=tokens, output_hidden_states=True)
        ref = {"last": o.last_hidden_state, "penultimate": hf.text_model.final_layer_norm(o.hidden_states[-2])}
    for layer, idx in (("last", 0), ("penultimate", 1)):
        z = text_encode(sd, tokens, heads, layers - idx)
        err = float((z - ref[layer]).norm() / ref[layer].norm())
        print(f"text_small[{layer}] restatement vs transformers.CLIPTextModel rel-L2 = {err:.3e}, |z| = {float(ref[layer].abs().mean()):.3f}")
        assert err < 1e-5
        out["z_" + layer] = _np(ref[layer])
    np.savez_compressed(os.path.join(GOLD, "text_small.npz"), **out)


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    which = sys.argv[1:] or ["tables", "ddim_step", "ddim_step_sam", "p_sample", "queue", "unet_small", "unet_small64", "unet_full", "vae_small", "text_small"]
    for w in which:
        print("==", w)
        globals()["gen_" + w]()

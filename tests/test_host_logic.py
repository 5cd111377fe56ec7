"""CPU tests of the host-side logic: schedule tables, FIFO geometry and the multi-rank window sharding (gloo, world 2/4).

The FIFO driver is exercised with a FAKE op namespace (torch CPU restatements built on the oracle) and a fake sampler,
so that the sharding / hand-off / RNG-order logic is tested without a GPU; the real kernels are tested under -m gpu."""
import os
import types

import numpy as np
import pytest
import torch

from conftest import GOLDEN


def test_schedule_tables_match_oracle_and_reference_golden():
    from moca_video_b200 import schedule
    from oracle import step_ref as R
    g = np.load(os.path.join(GOLDEN, "tables.npz"))
    for steps, eta in ((64, 1.0), (50, 1.0), (16, 0.0)):
        t = schedule.make_ddim_tables(steps, eta)
        o = R.make_tables(steps, eta)
        k = f"s{steps}_"
        assert np.array_equal(t.timesteps, g[k + "timesteps"])
        for n in ("alphas", "alphas_prev", "sigmas", "sqrt_one_minus_alphas", "scale", "scale_prev"):
            assert np.array_equal(getattr(t, n), g[k + n]), n
            assert np.array_equal(getattr(t, n), o[n]), n
        assert np.array_equal(t.coef([0, 3, 5]), R.coef_rows(o, [0, 3, 5]))


def test_fifo_geometry():
    from moca_video_b200 import schedule
    from moca_video_b200.fifo import window_plan
    from oracle import step_ref as R
    ts, idx = schedule.fifo_queue_tables(64, 16, True)
    ots, oidx = R.fifo_geometry(64, 16, True)
    assert np.array_equal(ts, ots) and np.array_equal(idx, oidx) and len(ts) == 72
    assert list(ts[:9]) == [0] * 9 and ts[-1] == 999 and ts[9] == 16
    fa = types.SimpleNamespace(video_length=16, num_partitions=4, lookahead_denoising=True)
    plan = window_plan(fa)
    assert [p[0] for p in plan] == [7, 6, 5, 4, 3, 2, 1, 0]
    assert plan[0][1:] == (56, 64, 72) and plan[-1][1:] == (0, 8, 16)
    assert schedule.reference_mask_index(16, 40) == [0] + [9] * 15
    fb = types.SimpleNamespace(video_length=16, num_partitions=4, lookahead_denoising=False)
    assert [p[1] for p in window_plan(fb)] == [48, 32, 16, 0]


# ---- fake backend ------------------------------------------------------------------------------------------------
def _fake_ops():
    from oracle import step_ref as R
    ns = types.SimpleNamespace()
    ns.window_gather = lambda q, start, n: q[:, :, start:start + n].clone()

    def scatter(q, win, off, start, n):
        q[:, :, start:start + n] = win[:, :, off:off + n]
        return q
    ns.window_scatter = scatter
    ns.queue_shift = lambda q, tail: torch.cat([q[:, :, 1:], tail], dim=2)
    ns.freeinit_filter = lambda shape, device, ft, n, ds, dt: torch.from_numpy(R.gaussian_low_pass_filter(*shape[-3:], ds, dt))
    ns.freq_mix_3d = lambda x, n, lpf: torch.from_numpy(R.freq_mix_3d(x.numpy(), n.numpy(), lpf.numpy()))
    return ns


class _FakeSampler:
    """x_prev depends on the window content, its timesteps and its noise: any sharding mistake changes the queue."""

    def fifo_onestep(self, cond, shape, latents, timesteps, indices, noise=None, **kw):
        t = torch.as_tensor(np.asarray(timesteps), dtype=torch.float32).view(1, 1, -1, 1, 1) / 1000.0
        mixed = latents.flip(2).cumsum(2).flip(2) * 0.05          # every frame sees the later frames of its window
        return 0.9 * latents + 0.02 * noise + 0.1 * t + mixed, latents


def _run(rank, world, port, out_path):
    import torch.distributed as dist
    from moca_video_b200 import fifo
    if world > 1:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    fifo.ops = _fake_ops()
    fa = types.SimpleNamespace(num_inference_steps=64, lookahead_denoising=True, video_length=16, num_partitions=4, new_video_length=4)
    g = torch.Generator().manual_seed(5)
    lat = torch.randn(1, 4, 72, 4, 6, generator=g)
    out = {}
    frames = fifo.fifo_ddim_sampling(fa, None, {"c_crossattn": []}, (1, 4, 16, 4, 6), _FakeSampler(), cfg_scale=1.0, latents=lat.clone(),
                                     generator=g, num_iterations=3, queue_out=out)
    if rank == 0:
        torch.save({"q": out["latents"], "frames": torch.cat(frames, 2)}, out_path)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _spawn(world, tmp_path, port):
    import torch.multiprocessing as mp
    p = str(tmp_path / f"w{world}.pt")
    if world == 1:
        _run(0, 1, port, p)
    else:
        mp.spawn(_run, args=(world, port, p), nprocs=world, join=True)
    return torch.load(p)


@pytest.mark.parametrize("world", [2, 4])
def test_fifo_sharding_matches_single_rank(world, tmp_path):
    """1-vs-G equality of the queue and of the popped frames after 3 iterations (same pre-drawn noise stream)."""
    import moca_video_b200.fifo as fifo_mod
    real_ops = fifo_mod.ops
    try:
        one = _spawn(1, tmp_path, 29611)
    finally:
        fifo_mod.ops = real_ops
    many = _spawn(world, tmp_path, 29611 + world)
    assert torch.equal(one["q"], many["q"])
    assert torch.equal(one["frames"], many["frames"])


def test_fifo_single_rank_matches_reference_order():
    """The in-place reversed window loop (funcs.py:306,351-352) equals the snapshot formulation the sharded path uses."""
    import moca_video_b200.fifo as fifo_mod
    real_ops = fifo_mod.ops
    try:
        fifo_mod.ops = _fake_ops()
        fa = types.SimpleNamespace(num_inference_steps=64, lookahead_denoising=True, video_length=16, num_partitions=4, new_video_length=4)
        g = torch.Generator().manual_seed(9)
        lat = torch.randn(1, 4, 72, 4, 6, generator=g)
        g2 = torch.Generator().manual_seed(10)
        out = {}
        fifo_mod.fifo_ddim_sampling(fa, None, {}, (1, 4, 16, 4, 6), _FakeSampler(), cfg_scale=1.0, latents=lat.clone(), generator=g2,
                                    num_iterations=1, queue_out=out)
        # snapshot restatement
        g2 = torch.Generator().manual_seed(10)
        noises = [torch.cat([torch.randn((1, 4, 1, 4, 6), generator=g2) for _ in range(16)], dim=2) for _ in range(8)]
        ts, idx = fifo_mod.fifo_queue_tables(64, 16, True)
        snap = lat.clone(); new = lat.clone()
        for (w, start, mid, end), nz in zip(fifo_mod.window_plan(fa), noises):
            xp, _ = _FakeSampler().fifo_onestep(None, None, snap[:, :, start:end], ts[start:end], idx[start:end], noise=nz)
            new[:, :, mid:end] = xp[:, :, 8:]
        tail = torch.randn((1, 4, 4, 6), generator=g2).unsqueeze(2)
        lpf = fifo_mod.ops.freeinit_filter((1, 4, 6), "cpu", "gaussian", 1, 0.25, 0.25)
        exp = torch.cat([new[:, :, 1:], fifo_mod.ops.freq_mix_3d(new[:, :, 0:1], tail, lpf)], dim=2)
        assert torch.allclose(out["latents"], exp, atol=1e-6)
    finally:
        fifo_mod.ops = real_ops
        fifo_mod._filter_cache.clear()

"""Diagnostic parity sweep for the GPU box: python tools/gpu_check.py [section ...]

Prints one line per case with the error against a plain PyTorch / oracle reference; never raises on a numerical
mismatch so that one run reports everything.  (The pass/fail versions of these checks live in tests/ under -m gpu.)
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from moca_video_b200 import ops  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = "cuda"


def rel(a, b):
    a = a.float(); b = b.float()
    return float((a - b).norm() / (b.norm() + 1e-30))


def sec_step():
    from oracle import step_ref as R
    g = np.load(os.path.join(ROOT, "tests/golden/step_ddim.npz"))
    tables = R.make_tables(64, 1.0)
    for name in "abc":
        x, e_c, e_u, noise = (g[f"{name}_{k}"] for k in ("x", "e_c", "e_u", "noise"))
        mask, cond, t, idx = g[f"{name}_mask"], g[f"{name}_cond"], g[f"{name}_t"], g[f"{name}_idx"]
        with_mask = int(g[f"{name}_with_mask"])
        coef = R.coef_rows(tables, idx)
        mi = R.parity_mask_index(16, x.shape[3]) if with_mask else None
        e = R.cfg_combine(e_c, e_u, float(g[f"{name}_cfg"]))
        xo, po = R.moca_ddim_step(x, e, noise, coef, t, mask if with_mask else None, mi, cond, 0.5)
        tt = lambda a: torch.from_numpy(a).to(dev)
        xp, p0 = ops.cfg_ddim_step(tt(x), tt(e_c), tt(e_u), tt(noise), float(g[f"{name}_cfg"]), coef, t,
                                   tt(mask) if with_mask else None, mi, tt(cond) if with_mask else None, 0.5)
        dx = np.abs(xp.cpu().numpy() - xo).max(); dp = np.abs(p0.cpu().numpy() - po).max()
        gx = np.abs(xp.cpu().numpy() - g[f"{name}_x_prev"]).max(); gp = np.abs(p0.cpu().numpy() - g[f"{name}_pred_x0"]).max()
        print(f"step[{name}] vs oracle: x_prev maxabs {dx:.3e} pred_x0 {dp:.3e} | vs reference golden: {gx:.3e} {gp:.3e}  bitexact={dx == 0 and dp == 0}")
    # p_sample
    g = np.load(os.path.join(ROOT, "tests/golden/step_psample.npz"))
    t50 = R.make_tables(50, 1.0)
    i = int(g["index"])
    for key, us in (("scale_", True), ("plain_", False)):
        tt = lambda a: torch.from_numpy(a).to(dev)
        xp, p0 = ops.p_sample_ddim(tt(g["x"]), tt(g["e_c"]), tt(g["e_u"]), tt(g[key + "noise"]), float(g["cfg"]), t50["alphas"][i],
                                   t50["alphas_prev"][i], t50["sigmas"][i], t50["sqrt_one_minus_alphas"][i],
                                   t50["scale"][i] if us else None, t50["scale_prev"][i] if us else None)
        print(f"p_sample[{key}] vs reference golden: x_prev {np.abs(xp.cpu().numpy() - g[key + 'x_prev']).max():.3e} "
              f"pred_x0 {np.abs(p0.cpu().numpy() - g[key + 'pred_x0']).max():.3e}")
    # queue
    g = np.load(os.path.join(ROOT, "tests/golden/queue_freeinit.npz"))
    t64 = R.make_tables(64, 1.0)
    z, qn = g["z"], g["queue_noise"]
    src = [0] * 8 + [max(0, i - (64 - z.shape[2])) for i in range(64)]
    al = [t64["alphas"][0]] * 8 + [t64["alphas"][i] for i in range(64)]
    q = ops.renoise(torch.from_numpy(z).to(dev), torch.from_numpy(qn).to(dev), src, al)
    print(f"renoise vs reference golden maxabs {np.abs(q.cpu().numpy() - g['queue']).max():.3e}")
    for ft in ("gaussian", "butterworth", "ideal", "box"):
        f = ops.freeinit_filter((8, 10, 12), dev, ft, 4, 0.25, 0.25)
        print(f"filter[{ft}] maxabs {np.abs(f.cpu().numpy() - g['filter_' + ft]).max():.3e}")
    f1 = ops.freeinit_filter((1, 40, 64), dev, "gaussian", 1, 0.25, 0.25)
    print(f"filter[gauss T=1] maxabs {np.abs(f1.cpu().numpy() - g['filter_gauss_t1']).max():.3e}")
    lpf = ops.freeinit_filter((8, 10, 12), dev, "gaussian", 1, 0.25, 0.25)
    y = ops.freq_mix_3d(torch.from_numpy(g["mix_x"]).to(dev), torch.from_numpy(g["mix_noise"]).to(dev), lpf)
    print(f"freq_mix_3d vs reference golden rel {rel(y.cpu(), torch.from_numpy(g['mix_out'])):.3e}")
    lat = torch.from_numpy(g["shift_in"]).to(dev)
    lp1 = ops.freeinit_filter((1, 6, 8), dev, "gaussian", 1, 0.25, 0.25)
    tail = ops.freq_mix_3d(lat[:, :, 0:1].contiguous(), torch.from_numpy(g["shift_noise"]).to(dev).reshape(1, 4, 1, 6, 8), lp1)
    sh = ops.queue_shift(lat, tail)
    print(f"shift_latents vs reference golden rel {rel(sh.cpu(), torch.from_numpy(g['shift_out'])):.3e}")
    x = torch.randn(8, 4, 16, 40, 64, device=dev); n = torch.randn_like(x)
    lpf = ops.freeinit_filter((16, 40, 64), dev, "gaussian", 1, 0.25, 0.25)
    y = ops.freq_mix_3d(x, n, lpf)
    ax = (-3, -2, -1)
    xf = torch.fft.fftshift(torch.fft.fftn(x, dim=ax), dim=ax); nf = torch.fft.fftshift(torch.fft.fftn(n, dim=ax), dim=ax)
    yr = torch.fft.ifftn(torch.fft.ifftshift(xf * lpf + nf * (1 - lpf), dim=ax), dim=ax).real
    print(f"freq_mix_3d [8,4,16,40,64] vs torch.fft rel {rel(y, yr):.3e}")


def sec_gemm():
    g = torch.Generator(device=dev).manual_seed(0)
    for (M, N, K) in [(128, 128, 64), (128, 128, 256), (256, 256, 128), (1000, 320, 320), (77, 2560, 1024), (16, 1280, 320),
                      (2560, 3840, 1280), (40960, 320, 320), (3000, 640, 1920), (5000, 4, 320), (300, 160, 64), (40960, 960, 320)]:
        a = torch.randn(M, K, device=dev, generator=g).to(ops.op_dtype())
        w = (torch.randn(N, K, device=dev, generator=g) / K ** 0.5).to(ops.op_dtype())
        b = torch.randn(N, device=dev, generator=g)
        ref = a.float() @ w.float().t() + b
        try:
            y = ops.gemm_op16(a, w, b, False)
            torch.cuda.synchronize()
            yb = ops.gemm_op16(a, w, None, True)
            torch.cuda.synchronize()
            print(f"gemm M{M} N{N} K{K}: fp32-out rel {rel(y, ref):.3e}  bf16-out rel {rel(yb, ref - b):.3e}")
        except Exception as e:  # noqa: BLE001
            print(f"gemm M{M} N{N} K{K}: EXC {e}")
            return
    M, K, N = 3000, 320, 2560
    a = torch.randn(M, K, device=dev, generator=g).to(ops.op_dtype())
    w = torch.randn(N, K, device=dev, generator=g) / K ** 0.5
    b = torch.randn(N, device=dev, generator=g)
    h = a.float() @ w.to(ops.op_dtype()).float().t() + b
    v, gate = h.chunk(2, dim=-1)
    ref = v * torch.nn.functional.gelu(gate)
    y = ops.geglu(a, w, b)
    print(f"geglu M{M} N{N} K{K}: rel {rel(y, ref):.3e}")


def sec_conv():
    import torch.nn.functional as F
    g = torch.Generator(device=dev).manual_seed(1)
    for (NF, H, W, Ci, Co, s) in [(2, 8, 8, 64, 64, 1), (3, 10, 16, 64, 128, 1), (16, 40, 64, 320, 320, 1), (4, 20, 32, 960, 640, 1),
                                  (2, 5, 8, 1280, 1280, 1), (4, 40, 64, 320, 320, 2), (16, 10, 16, 64, 4, 1), (2, 8, 8, 8, 64, 1)]:
        x = torch.randn(NF, H, W, Ci, device=dev, generator=g)
        w = torch.randn(Co, Ci, 3, 3, device=dev, generator=g) / (9 * Ci) ** 0.5
        b = torch.randn(Co, device=dev, generator=g)
        ref = F.conv2d(x.to(ops.op_dtype()).float().permute(0, 3, 1, 2), w.to(ops.op_dtype()).float(), b, stride=s, padding=1).permute(0, 2, 3, 1)
        try:
            y = ops.conv3x3_nhwc(x, w, b, s)
            print(f"conv3x3 NF{NF} {H}x{W} {Ci}->{Co} s{s}: rel {rel(y, ref):.3e}")
        except Exception as e:  # noqa: BLE001
            print(f"conv3x3 NF{NF} {H}x{W} {Ci}->{Co} s{s}: EXC {e}")
            return
    for (B, T, HW, C) in [(1, 16, 40, 64), (2, 4, 100, 128), (1, 16, 2560, 320), (2, 16, 160, 1280)]:
        x = torch.randn(B, T, HW, C, device=dev, generator=g)
        w = torch.randn(C, C, 3, 1, 1, device=dev, generator=g) / (3 * C) ** 0.5
        b = torch.randn(C, device=dev, generator=g)
        x5 = x.to(ops.op_dtype()).float().permute(0, 3, 1, 2).reshape(B, C, T, HW, 1)
        ref = F.conv3d(x5, w.to(ops.op_dtype()).float(), b, padding=(1, 0, 0)).reshape(B, C, T, HW).permute(0, 2, 3, 1)
        y = ops.tconv_tokens(x, w, b)
        print(f"tconv B{B} T{T} HW{HW} C{C}: rel {rel(y, ref):.3e}")


def sec_norm():
    import torch.nn.functional as F
    g = torch.Generator(device=dev).manual_seed(2)
    for (NF, HW, C, rps_frames, eps, silu) in [(16, 2560, 320, 1, 1e-5, 1), (16, 160, 1280, 16, 1e-6, 0), (4, 640, 960, 1, 1e-5, 1),
                                               (2, 40, 2560, 2, 1e-5, 1), (8, 64, 64, 1, 1e-5, 1)]:
        x = torch.randn(NF * HW, C, device=dev, generator=g) * 2 + 0.5
        ga = torch.randn(C, device=dev, generator=g); be = torch.randn(C, device=dev, generator=g)
        y = ops.groupnorm_op16(x, ga, be, HW * rps_frames, eps, silu)
        xr = x.reshape(NF // rps_frames, rps_frames * HW, C).permute(0, 2, 1)
        ref = F.group_norm(xr, 32, ga, be, eps)
        if silu:
            ref = F.silu(ref)
        ref = ref.permute(0, 2, 1).reshape(NF * HW, C)
        print(f"groupnorm rows{NF * HW} C{C} span{rps_frames} silu{silu}: rel {rel(y, ref):.3e}")
    for (rows, C) in [(1000, 320), (333, 1280), (64, 64), (4096, 640), (100, 512)]:
        x = torch.randn(rows, C, device=dev, generator=g) * 3 + 1
        ga = torch.randn(C, device=dev, generator=g); be = torch.randn(C, device=dev, generator=g)
        y = ops.layernorm_op16(x, ga, be)
        print(f"layernorm rows{rows} C{C}: rel {rel(y, F.layer_norm(x, (C,), ga, be, 1e-5)):.3e}")


def sec_attn():
    g = torch.Generator(device=dev).manual_seed(3)

    def ref_attn(q, k, v, batch, heads, Lq, Lk, Lv):
        q = q.float().reshape(batch, Lq, heads, 64).permute(0, 2, 1, 3)
        k = k.float().reshape(batch, Lk, heads, 64).permute(0, 2, 1, 3)[:, :, :Lv]
        v = v.float().reshape(batch, Lk, heads, 64).permute(0, 2, 1, 3)[:, :, :Lv]
        s = (q @ k.transpose(-1, -2)) * 0.125
        return (s.softmax(-1) @ v).permute(0, 2, 1, 3).reshape(batch * Lq, heads * 64)

    for (batch, heads, Lq, Lk, Lv) in [(2, 1, 64, 64, 64), (3, 5, 40, 40, 40), (2, 20, 160, 160, 160), (4, 10, 640, 640, 640),
                                       (2, 5, 2560, 2560, 2560), (4, 5, 640, 154, 77), (2, 20, 40, 77, 77), (2, 5, 100, 154, 154)]:
        q = torch.randn(batch * Lq, heads * 64, device=dev, generator=g).to(ops.op_dtype())
        k = torch.randn(batch * Lk, heads * 64, device=dev, generator=g).to(ops.op_dtype())
        v = torch.randn(batch * Lk, heads * 64, device=dev, generator=g).to(ops.op_dtype())
        try:
            y = ops.attention_op16(q, k, v, batch, heads, Lq, Lk, Lv)
            print(f"attention b{batch} h{heads} Lq{Lq} Lk{Lk} valid{Lv}: rel {rel(y, ref_attn(q, k, v, batch, heads, Lq, Lk, Lv)):.3e}")
        except Exception as e:  # noqa: BLE001
            print(f"attention b{batch} h{heads} Lq{Lq} Lk{Lk}: EXC {e}")
            return
    # temporal: rows ordered (b, t, hw); sequence o = (b, hw)
    for (B, T, HW, heads) in [(1, 16, 40, 5), (2, 16, 160, 20), (1, 4, 64, 1), (1, 16, 2560, 5)]:
        C = heads * 64
        qkv = torch.randn(B * T * HW, 3 * C, device=dev, generator=g).to(ops.op_dtype())
        q, k, v = qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:]
        y = ops.attention_op16(q, k, v, B * HW, heads, T, T, T, temporal_inner=HW)
        perm = lambda t: t.float().reshape(B, T, HW, C).permute(0, 2, 1, 3).reshape(B * HW * T, C)
        r = ref_attn(perm(q), perm(k), perm(v), B * HW, heads, T, T, T)
        r = r.reshape(B, HW, T, C).permute(0, 2, 1, 3).reshape(B * T * HW, C)
        print(f"temporal attention B{B} T{T} HW{HW} h{heads}: rel {rel(y, r):.3e}")


def _unet_case(cfg, b, t, h, w, L, fifo, seed, taps=True, ctx_len=None):
    from moca_video_b200.arch import random_state_dict
    from moca_video_b200.unet import UNet3D
    from oracle.unet_ref import unet_forward
    sd = random_state_dict(cfg, seed=0, device=dev)
    net = UNet3D(cfg.in_channels, cfg.out_channels, cfg.model_channels, cfg.attention_resolutions, cfg.num_res_blocks,
                 cfg.channel_mult, cfg.num_head_channels, cfg.context_dim, cfg.init_attn_heads, device=dev)
    net.load_state_dict(sd)
    g = torch.Generator(device=dev).manual_seed(seed)
    x = torch.randn(b, cfg.in_channels, t, h, w, device=dev, generator=g)
    ctx = torch.randn(b, L, cfg.context_dim, device=dev, generator=g)
    ts = torch.randint(0, 1000, (t if fifo else b,), device=dev, generator=g)
    rt = {}
    t0 = time.time()
    y_ref = unet_forward(sd, cfg, x, ts, ctx, fps=10, taps=rt)
    torch.cuda.synchronize()
    t_ref = time.time() - t0
    if taps:
        net.set_taps(True)
        y = net(x, ts, ctx, fps=10)
        torch.cuda.synchronize()
        mt = net.taps()
        for k, v in mt.items():
            if k in rt:
                print(f"   tap {k:32s} rel {rel(v, rt[k]):.3e}  |ref| {float(rt[k].abs().mean()):.3f}")
        print(f"  taps-mode output rel {rel(y, y_ref):.3e}")
        net.set_taps(False)
    y = net(x, ts, ctx, fps=10)
    torch.cuda.synchronize()
    print(f"  graph-mode output rel-L2 {rel(y, y_ref):.3e}  |y_ref| {float(y_ref.abs().mean()):.4f}  torch-fp32 oracle {t_ref * 1e3:.1f} ms")
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        net(x, ts, ctx, fps=10)
    ev0.record()
    n = 10
    for _ in range(n):
        net(x, ts, ctx, fps=10)
    ev1.record(); torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / n
    print(f"  forward {ms:.3f} ms  launches {net.launches()}  flops {net.flops() / 1e12:.3f} T  -> {net.flops() / ms / 1e9:.1f} TFLOP/s  "
          f"workspace {net.workspace_bytes() / 1e9:.2f} GB weights {net.weight_bytes() / 1e9:.2f} GB")
    return net


def sec_unet_small():
    from moca_video_b200.arch import UNetConfig
    cfg = UNetConfig(model_channels=64, context_dim=64)
    print("unet small (mc=64) fifo b1 t16 8x8 L11")
    _unet_case(cfg, 1, 16, 8, 8, 11, True, 5)
    print("unet small (mc=64) plain b2 t4 8x16 L7")
    _unet_case(cfg, 2, 4, 8, 16, 7, False, 6)


def sec_unet_full():
    from moca_video_b200.arch import UNetConfig
    cfg = UNetConfig.videocrafter2()
    print("unet full b1 t16 40x64 L77 (fifo timesteps)")
    _unet_case(cfg, 1, 16, 40, 64, 77, True, 7, taps=True)


def sec_unet_full_b2():
    from moca_video_b200.arch import UNetConfig
    cfg = UNetConfig.videocrafter2()
    print("unet full b2 t16 40x64 L77 (CFG batch)")
    _unet_case(cfg, 2, 16, 40, 64, 77, False, 8, taps=False)


if __name__ == "__main__":
    secs = sys.argv[1:] or ["step", "gemm", "conv", "norm", "attn", "unet_small", "unet_full"]
    for s in secs:
        print(f"=== {s}", flush=True)
        try:
            globals()["sec_" + s]()
        except Exception as e:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            print(f"=== {s} FAILED: {e}")
        sys.stdout.flush()

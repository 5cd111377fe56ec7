"""numpy restatement of the sampler / queue / FreeInit arithmetic of the MoCA-Video hot path.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Every function cites the reference
file:line it follows (paths relative to ``/root/reference``).  All arithmetic is float32 in the
reference's operation order (torch eager does not contract mul+add into FMA), so the CUDA step
kernels are expected to match these functions bit for bit.

Pinned by ``tests/golden/step_*.npz`` which were produced by the unmodified reference
(``oracle/make_golden.py``).
"""
import math

import numpy as np

F32 = np.float32


# --------------------------------------------------------------------------------------------
# schedules (host side tables)
# --------------------------------------------------------------------------------------------
def make_beta_schedule_linear(n=1000, linear_start=0.00085, linear_end=0.012):
    """lvdm/models/utils_diffusion.py:31-36 ("linear") with the yaml's range (configs/...yaml:4-7)."""
    return np.linspace(linear_start ** 0.5, linear_end ** 0.5, n, dtype=np.float64) ** 2


def alphas_cumprod_f32(n=1000):
    """lvdm/models/ddpm3d.py:113-134: float64 cumprod registered as a float32 buffer."""
    betas = make_beta_schedule_linear(n)
    return np.cumprod(1.0 - betas, axis=0).astype(F32)


def make_ddim_timesteps(num_ddim, num_ddpm=1000):
    """lvdm/models/utils_diffusion.py:56-66 ('uniform' = linspace().round() variant)."""
    return np.linspace(0, num_ddpm - 1, num_ddim).round().astype(np.int64)


def scale_arr_f32():
    """lvdm/models/ddpm3d.py:362-376 (the kept "bug": 400 + 1000 entries)."""
    return np.concatenate((np.linspace(1.0, 0.7, 400), np.full(1000, 0.7))).astype(F32)


def make_tables(num_ddim, eta, num_ddpm=1000):
    """DDIMSampler.make_schedule, lvdm/models/samplers/ddim.py:62-106 + utils_diffusion.py:81-93.

    Mixed dtypes of the reference are reproduced and every table is finally cast to
    float32, which is what ``torch.full(size, table[index])`` does at the point of use
    (ddim.py:409-412).  Returns dict of float32 arrays of length ``num_ddim`` plus ``timesteps``.
    """
    ac = alphas_cumprod_f32(num_ddpm)                      # float32 tensor in the reference
    ts = make_ddim_timesteps(num_ddim, num_ddpm)
    alphas = ac[ts]                                         # float32
    alphas_prev = np.asarray([float(ac[0])] + ac[ts[:-1]].astype(np.float64).tolist())  # float64 ndarray
    # ndarray / Tensor dispatches to Tensor.__rtruediv__ == reciprocal(tensor) * ndarray: the reciprocal of
    # (1 - alphas) is taken in float32 and only then promoted to float64 (utils_diffusion.py:88)
    recip = (F32(1.0) / (F32(1.0) - alphas)).astype(np.float64)
    sigmas = eta * np.sqrt(recip * (1 - alphas_prev) * (1 - alphas.astype(np.float64) / alphas_prev))
    sqrt_1m = np.sqrt(F32(1.0) - alphas).astype(F32)        # np.sqrt(1. - ddim_alphas) on a float32 tensor
    sc = scale_arr_f32()
    scale = sc[ts]                                          # ddim.py:78-79
    scale_prev = np.asarray([sc[0]] + sc[ts[:-1]].tolist()).astype(F32)  # ddim.py:80-81
    return dict(timesteps=ts, alphas=alphas.astype(F32), alphas_prev=alphas_prev.astype(F32),
                sigmas=sigmas.astype(F32), sqrt_one_minus_alphas=sqrt_1m,
                scale=scale.astype(F32), scale_prev=scale_prev)


def fifo_geometry(num_steps=64, video_length=16, lookahead=True, num_ddpm=1000):
    """scripts/evaluation/funcs.py:290-294: queue timesteps / table indices with look-ahead copies."""
    ts = make_ddim_timesteps(num_steps, num_ddpm)
    idx = np.arange(num_steps)
    if lookahead:
        ts = np.concatenate([np.full((video_length // 2,), ts[0]), ts])
        idx = np.concatenate([np.full((video_length // 2,), 0), idx])
    return ts, idx


def coef_rows(tables, indices):
    """The float[4][F] coefficient block the C-ABI takes: A, A_prev, sigma, sqrt(1-A) per frame."""
    indices = np.asarray(indices)
    return np.stack([tables["alphas"][indices], tables["alphas_prev"][indices],
                     tables["sigmas"][indices], tables["sqrt_one_minus_alphas"][indices]]).astype(F32)


def parity_mask_index(F, H):
    """the frame loop variable is clobbered by the visualisation loops
    (ddim.py:405,477,502,533) so frames >= 1 read mask ceil(H/4)-1."""
    j = math.ceil(H / 4) - 1
    return np.asarray([0] + [j] * (F - 1), dtype=np.int32)


# --------------------------------------------------------------------------------------------
# K13: CFG + MoCA ddim_step
# --------------------------------------------------------------------------------------------
def cfg_combine(e_c, e_u, scale):
    """ddim.py:372  e = e_u + s * (e_c - e_u); skipped when uc is None or s == 1 (ddim.py:366)."""
    if e_u is None or scale == 1.0:
        return e_c.astype(F32)
    return (e_u + F32(scale) * (e_c - e_u)).astype(F32)


def moca_ddim_step(x, e, noise, coef, ts, mask=None, mask_idx=None, cond_img=None, gamma=0.5,
                   enh_lo=1.5, enh_hi=1.0, t_thresh=300, cover_max=0.0, sqrt=np.sqrt):
    """DDIMSampler.ddim_step, ddim.py:377-649 with SURVEY Q1-Q5 resolved (Appendix B).

    x, e, noise: [B,C,F,H,W] float32 (``noise`` = the N(0,1) draws of ddim.py:561, frame order).
    coef: float32 [4,F] (A, A_prev, sigma, sqrt(1-A));  ts: int [F].
    mask: [B,1,Fm,H,W] or None; mask_idx[f] = mask frame used by frame f (-1: no blend).
    cond_img: [B,C,1,H,W].  enh_hi == 0 disables the blend for t > t_thresh (SAM branch :592);
    cover_max: skip masks whose sum exceeds cover_max*numel; <= 0 (default) = no coverage test, which is what the DAVIS
    branch does (ddim.py:587); the SAM branch's 80 % rule is ``sam_blend_frame``.
    sqrt: scalar float32 square root.  Default = IEEE correctly rounded (numpy; CUDA ``sqrtf`` is the same,
    and the reference runs on CUDA).  torch's *CPU* sqrt is 1 ulp off for some inputs, so the test that
    compares with the CPU-generated goldens passes ``torch_cpu_sqrt`` to get bit equality.
    Returns x_prev, pred_x0.
    """
    x = x.astype(F32); e = e.astype(F32); noise = noise.astype(F32)
    B, C, Fr, H, W = x.shape
    x_prev = np.empty_like(x); pred = np.empty_like(x)
    beta = F32(0.9); omb = F32(1 - 0.9)
    m_prev = np.zeros((B, C, H, W), F32)
    p_prev = None
    for i in range(Fr):
        A, Ap, Sg, R = (F32(coef[0, i]), F32(coef[1, i]), F32(coef[2, i]), F32(coef[3, i]))
        xi, ei = x[:, :, i], e[:, :, i]
        p = (xi - R * ei) / sqrt(A)                                      # :415
        cdir = sqrt((F32(1.0) - Ap) - Sg * Sg)                           # :418
        d = cdir * ei
        if p_prev is not None:
            g = (p - p_prev) + F32(1.5) * d                              # :422-423
            m = beta * m_prev + omb * g                                  # :424-427
            corr = F32(2) * (F32(1.0) - F32(ts[i]) / F32(1000.0))        # :428
            p = p + corr * m                                             # :430,557
            m_prev = m
        p_prev = p                                                       # :559
        n = Sg * noise[:, :, i]                                          # :561
        x_prev[:, :, i] = (sqrt(Ap) * p + d) + n                         # :562
        if mask is not None and mask_idx is not None and mask_idx[i] >= 0 and mask_idx[i] < mask.shape[2]:
            k = enh_lo if ts[i] <= t_thresh else enh_hi                  # :584 / :592,:845
            mk = mask[:, :, mask_idx[i]].astype(F32)                     # [B,1,H,W]
            s = float(mk.astype(np.float64).sum())
            if k != 0 and s != 0 and (cover_max <= 0 or s <= cover_max * mk.size):     # :587
                ci = cond_img[:, :, 0].astype(F32) * F32(k)
                p = np.where(mk > 0.5, ci, p)                            # :588-590
        pred[:, :, i] = F32(1 - gamma) * p + F32(gamma) * n              # :609
    return x_prev, pred


# --------------------------------------------------------------------------------------------
# SAM branch of ddim_step (no DAVIS masks): segmentation glue + blend
# --------------------------------------------------------------------------------------------
def latent_frame_to_u8(frame):
    """DDIMSampler._apply_segmentation, ddim.py:745-761: the [C,H,W] x0 prediction of one frame -> the uint8 RGB image
    handed to Grounding-DINO / SAM-2.  First three channels (a single channel is repeated); values are scaled by 255
    when floor(max) <= 1, else taken as they are; ``astype(np.uint8)`` of a float truncates toward zero and wraps
    modulo 256 (what numpy does on x86-64 for values inside the int32 range; the golden pins it)."""
    f = np.transpose(np.asarray(frame, F32), (1, 2, 0))
    if f.shape[-1] != 3:
        f = np.repeat(f, 3, axis=-1) if f.shape[-1] == 1 else f[:, :, :3]
    if np.floor(f.max()) <= 1.0:
        f = f * F32(255)
    return (np.trunc(f).astype(np.int64) & 255).astype(np.uint8)


def mask_iou(m1, m2):
    """DDIMSampler.calculate_iou, ddim.py:905-943: mean IoU over zip(masks1, masks2) of the masks binarised at 0.5
    (an empty pair list gives nan, which never triggers the ``iou < 0.5`` fallback)."""
    ious = []
    for a, b in zip(np.asarray(m1) > 0.5, np.asarray(m2) > 0.5):
        inter = F32(np.logical_and(a, b).sum())
        union = F32(np.logical_or(a, b).sum())
        ious.append(F32(1.0) if union == 0 else inter / union)
    return float(np.mean(np.asarray(ious, F32), dtype=F32)) if ious else float("nan")


def sam_gate_masks(new_masks, pre_masks):
    """ddim.py:787-813: masks to blend with for this frame.  ``new_masks`` None / empty = the detector found no box:
    fall back to the previous frame's masks (None: no blend at all); otherwise the new masks unless their mean IoU
    with the previous ones is below 0.5."""
    if new_masks is None or len(new_masks) == 0:
        return pre_masks
    if pre_masks is not None and mask_iou(new_masks, pre_masks) < 0.5:
        return pre_masks
    return np.asarray(new_masks, F32)


def sam_blend_frame(p, cond_img, masks, gain=2.0, cover_max=0.8):
    """ddim.py:818-901 for one frame.  p [B,C,H,W]; cond_img [B,C,H,W]; masks [N,H,W].  Masks are applied in order;
    a mask covering more than ``cover_max`` of the frame RESETS the result to the unblended prediction (:826-828) and
    is otherwise skipped.  (As shipped the reference's ``torch.where`` broadcasts [1,C,H,W] against [1,C,1,H,W] and
    returns C identical copies along dim 2; this restatement returns one of them, SURVEY Q2: every caller drops it.)"""
    out = p
    for mk in np.asarray(masks, F32):
        if float(mk.astype(np.float64).sum()) > cover_max * mk.size:
            out = p
            continue
        out = np.where(mk[None, None] > 0.5, cond_img.astype(F32) * F32(gain), out)
    return out.astype(F32)


def moca_ddim_step_sam(x, e, noise, coef, ts, cond_img, segment_fn, gamma=0.5, t_thresh=300, sqrt=np.sqrt):
    """DDIMSampler.ddim_step without DAVIS masks (ddim.py:377-430, 557-562, 591-609): frames with t <= 300 go through
    apply_cond_img -> _apply_segmentation (:739-903).  ``segment_fn(frame_index, image_u8[H,W,3])`` stands in for
    Grounding-DINO + SAM-2 and returns masks [N,H,W] or None (no box).  Returns x_prev, pred_x0, the per-frame mask
    lists actually blended and the uint8 images shown to the segmenter."""
    x = x.astype(F32); e = e.astype(F32); noise = noise.astype(F32)
    B, C, Fr, H, W = x.shape
    # momentum-corrected predictions and x_prev are what moca_ddim_step computes without masks (gamma = 0 returns p')
    x_prev, p_corr = moca_ddim_step(x, e, noise, coef, ts, gamma=0.0, sqrt=sqrt)
    pred = np.empty_like(x)
    pre = None
    used, images = [], []
    for i in range(Fr):
        p = p_corr[:, :, i]
        if ts[i] <= t_thresh:
            img = latent_frame_to_u8(p[0])
            images.append(img)
            masks = sam_gate_masks(segment_fn(i, img), pre)
            if masks is not None:
                p = sam_blend_frame(p, cond_img[:, :, 0], masks)
            pre = masks
            used.append(masks)
        else:
            used.append(None)
        n = F32(coef[2, i]) * noise[:, :, i]
        pred[:, :, i] = F32(1 - gamma) * p + F32(gamma) * n              # :609
    return x_prev, pred, used, images


def torch_cpu_sqrt(v):
    """float32 sqrt as torch's CPU kernel rounds it (test helper for the CPU-generated goldens)."""
    import torch
    return F32(torch.full((1,), float(v), dtype=torch.float32).sqrt().item())


def p_sample_ddim(x, e, noise, A, Ap, Sg, R, scale=None, scale_prev=None, sqrt=np.sqrt):
    """DDIMSampler.p_sample_ddim, ddim.py:339-357 (scalar index; ``use_scale`` branch :349-355)."""
    x = x.astype(F32); e = e.astype(F32)
    A, Ap, Sg, R = F32(A), F32(Ap), F32(Sg), F32(R)
    p = (x - R * e) / sqrt(A)
    d = sqrt((F32(1.0) - Ap) - Sg * Sg) * e
    n = Sg * noise.astype(F32)
    if scale is not None:
        p = p / F32(scale)
        xp = ((sqrt(Ap) * F32(scale_prev)) * p + d) + n
    else:
        xp = (sqrt(Ap) * p + d) + n
    return xp.astype(F32), p.astype(F32)


# --------------------------------------------------------------------------------------------
# K15: queue construction / shift
# --------------------------------------------------------------------------------------------
def prepare_latents(z, noise, alphas, video_length=16, num_steps=64, lookahead=True):
    """scripts/evaluation/funcs.py:55-80.  z [B,C,Fz,H,W] clean latents; noise [B,C,Q,H,W] pre-drawn in
    call order (look-ahead copies first); alphas = ddim_alphas (float32, len num_steps)."""
    z = z.astype(F32)
    out = []
    q = 0
    if lookahead:
        a = F32(alphas[0])
        for _ in range(video_length // 2):
            out.append(a ** F32(0.5) * z[:, :, [0]] + (F32(1) - a) ** F32(0.5) * noise[:, :, [q]]); q += 1
    for i in range(num_steps):
        a = F32(alphas[i])
        fi = max(0, i - (num_steps - z.shape[2]))
        out.append(a ** F32(0.5) * z[:, :, [fi]] + (F32(1) - a) ** F32(0.5) * noise[:, :, [q]]); q += 1
    return np.concatenate(out, axis=2).astype(F32)


def gaussian_low_pass_filter(T, H, W, d_s=0.25, d_t=0.25):
    """utils/freeinit_utils.py:73-91 (python float math, stored to a float32 tensor)."""
    m = np.zeros((T, H, W), F32)
    if d_s == 0 or d_t == 0:
        return m
    for t in range(T):
        for h in range(H):
            for w in range(W):
                d2 = ((d_s / d_t) * (2 * t / T - 1)) ** 2 + (2 * h / H - 1) ** 2 + (2 * w / W - 1) ** 2
                m[t, h, w] = math.exp(-1 / (2 * d_s ** 2) * d2)
    return m


def butterworth_low_pass_filter(T, H, W, n=4, d_s=0.25, d_t=0.25):
    """utils/freeinit_utils.py:94-113."""
    m = np.zeros((T, H, W), F32)
    if d_s == 0 or d_t == 0:
        return m
    for t in range(T):
        for h in range(H):
            for w in range(W):
                d2 = ((d_s / d_t) * (2 * t / T - 1)) ** 2 + (2 * h / H - 1) ** 2 + (2 * w / W - 1) ** 2
                m[t, h, w] = 1 / (1 + (d2 / d_s ** 2) ** n)
    return m


def ideal_low_pass_filter(T, H, W, d_s=0.25, d_t=0.25):
    """utils/freeinit_utils.py:116-134."""
    m = np.zeros((T, H, W), F32)
    if d_s == 0 or d_t == 0:
        return m
    for t in range(T):
        for h in range(H):
            for w in range(W):
                d2 = ((d_s / d_t) * (2 * t / T - 1)) ** 2 + (2 * h / H - 1) ** 2 + (2 * w / W - 1) ** 2
                m[t, h, w] = 1 if d2 <= d_s * 2 else 0
    return m


def box_low_pass_filter(T, H, W, d_s=0.25, d_t=0.25):
    """utils/freeinit_utils.py:137-157."""
    m = np.zeros((T, H, W), F32)
    if d_s == 0 or d_t == 0:
        return m
    thr_s = round(int(H // 2) * d_s)
    thr_t = round(T // 2 * d_t)
    cf, cr, cc = T // 2, H // 2, W // 2
    m[cf - thr_t:cf + thr_t, cr - thr_s:cr + thr_s, cc - thr_s:cc + thr_s] = 1.0
    return m


def freq_mix_3d(x, noise, lpf):
    """utils/freeinit_utils.py:7-47.  x, noise: [..., T, H, W]; lpf broadcastable [T,H,W]."""
    ax = (-3, -2, -1)
    xf = np.fft.fftshift(np.fft.fftn(x.astype(np.float64), axes=ax), axes=ax)
    nf = np.fft.fftshift(np.fft.fftn(noise.astype(np.float64), axes=ax), axes=ax)
    mixed = xf * lpf + nf * (1 - lpf.astype(np.float64))
    return np.fft.ifftn(np.fft.ifftshift(mixed, axes=ax), axes=ax).real.astype(F32)


def shift_latents(latents, new_noise, d_s=0.25, d_t=0.25):
    """scripts/evaluation/funcs.py:86-99 (non-DAVIS branch): anchor = frame 0, shift left by one,
    tail = freq_mix_3d(anchor, new_noise, gaussian LPF for T=1)."""
    B, C, Q, H, W = latents.shape
    anchor = latents[:, :, 0:1].copy()
    if new_noise.ndim == 4:                       # randn_like(latents[:, :, -1]).unsqueeze(2), funcs.py:93
        new_noise = new_noise[:, :, None]
    out = latents.copy()
    out[:, :, :-1] = latents[:, :, 1:]
    lpf = gaussian_low_pass_filter(1, H, W, d_s, d_t)
    out[:, :, -1:] = freq_mix_3d(anchor, new_noise, lpf)
    return out


def fifo_iteration(latents, eps_fn, noise, tables, cfg_scale, num_steps=64, video_length=16, n_part=4,
                   mask=None, cond_img=None, gamma=0.5):
    """One pass of HOT LOOP 2, scripts/evaluation/funcs.py:306-355 with look-ahead denoising:
    windows r = 2n-1 .. 0 of 16 frames at stride 8; only the later half is written back (:351-352).
    eps_fn(x[B,C,16,H,W], t[16]) -> (e_c, e_u).  noise: [2n, B,C,16,H,W] indexed by rank."""
    ts_q, idx_q = fifo_geometry(num_steps, video_length, True)
    half = video_length // 2
    lat = latents.copy()
    for r in reversed(range(2 * n_part)):
        s = r * half
        t = ts_q[s:s + video_length]; idx = idx_q[s:s + video_length]
        win = lat[:, :, s:s + video_length].copy()
        e_c, e_u = eps_fn(win, t)
        e = cfg_combine(e_c, e_u, cfg_scale)
        mi = parity_mask_index(video_length, win.shape[3]) if mask is not None else None
        xp, _ = moca_ddim_step(win, e, noise[r], coef_rows(tables, idx), t,
                               mask[:, :, s:s + video_length] if mask is not None else None, mi, cond_img, gamma)
        lat[:, :, s + half:s + video_length] = xp[:, :, -half:]
    return lat

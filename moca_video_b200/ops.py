"""Tensor-level wrappers over the C-ABI (torch tensors in, torch tensors out; torch only owns the memory).

Each function states the reference method it stands in for (paths relative to the reference tree).  Inputs must be
CUDA fp32 contiguous tensors in the reference's own layout; nothing is computed on the host.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check, ctx, host_f32, host_i32, lib, ptr, stream_ptr


def op_dtype():
    """torch dtype of the library's 16-bit operand buffers (moca_operand_dtype: fp16 by default, bf16 with -DMOCA_OPERAND_BF16)."""
    return torch.bfloat16 if lib().moca_operand_dtype() == 1 else torch.float16


def _f32c(t, name):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise _lib.MocaError(f"{name} must be a CUDA tensor (no CPU path exists)")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def cfg_ddim_step(x, e_c, e_u, noise, cfg_scale, coef, ts, mask=None, mask_idx=None, cond_img=None, gamma=0.5,
                  enh_lo=1.5, enh_hi=1.0, t_thresh=300, cover_max=0.0):
    """DDIMSampler.unet CFG combine + DDIMSampler.ddim_step (lvdm/models/samplers/ddim.py:372, 377-649).

    x, e_c, e_u, noise: [B,C,F,H,W]; coef: float32 [4,F] (A, A_prev, sigma, sqrt(1-A)); ts: int [F].
    Returns (x_prev, pred_x0)."""
    x = _f32c(x, "x"); e_c = _f32c(e_c, "e_c"); noise = _f32c(noise, "noise")
    e_u = _f32c(e_u, "e_u") if e_u is not None else None
    B, Cc, F, H, W = x.shape
    coef = np.ascontiguousarray(np.asarray(coef, dtype=np.float32))
    assert coef.shape == (4, F), coef.shape
    xp = torch.empty_like(x); p0 = torch.empty_like(x)
    Fm = 0
    if mask is not None:
        mask = _f32c(mask, "mask"); cond_img = _f32c(cond_img, "cond_img")
        Fm = mask.shape[2]
        assert mask.shape == (B, 1, Fm, H, W) and cond_img.shape == (B, Cc, 1, H, W)
        midx = host_i32(mask_idx)
    else:
        midx = None
    ts_h = host_i32(ts)
    with torch.cuda.device(x.device):
        check(lib().moca_cfg_ddim_step(ctx(x.device.index), ptr(x), ptr(e_c), ptr(e_u), ptr(noise), float(cfg_scale),
                                       coef.ctypes.data_as(C.c_void_p), C.cast(ts_h, C.c_void_p), ptr(mask), Fm,
                                       C.cast(midx, C.c_void_p) if midx is not None else None, ptr(cond_img), float(gamma),
                                       float(enh_lo), float(enh_hi), int(t_thresh), float(cover_max), B, Cc, F, H, W,
                                       ptr(xp), ptr(p0), stream_ptr()))
    return xp, p0


def latent_to_u8(pred_x0, frame):
    """_apply_segmentation's image conversion (ddim.py:745-761): frame ``frame`` of sample 0 of pred_x0 [B,C,F,H,W] ->
    uint8 [H,W,3] (device tensor) for Grounding-DINO / SAM-2."""
    pred_x0 = _f32c(pred_x0, "pred_x0")
    B, Cc, F, H, W = pred_x0.shape
    img = torch.empty((H, W, 3), device=pred_x0.device, dtype=torch.uint8)
    with torch.cuda.device(pred_x0.device):
        check(lib().moca_latent_to_u8(ctx(pred_x0.device.index), ptr(pred_x0), Cc, F, H, W, int(frame), ptr(img), stream_ptr()))
    return img


def mask_iou(masks_a, masks_b):
    """DDIMSampler.calculate_iou (ddim.py:905-943): mean IoU over zip(masks_a, masks_b) binarised at 0.5 -> python float."""
    if masks_a.numel() == 0 or masks_b.numel() == 0:
        return float("nan")                                  # torch.tensor([]).mean(): never below the 0.5 gate
    a = _f32c(masks_a, "masks_a"); b = _f32c(masks_b, "masks_b")
    H, W = a.shape[-2:]
    assert b.shape[-2:] == (H, W)
    out = C.c_float()
    with torch.cuda.device(a.device):
        check(lib().moca_mask_iou(ctx(a.device.index), ptr(a), a.reshape(-1, H, W).shape[0], ptr(b), b.reshape(-1, H, W).shape[0],
                                  H, W, C.byref(out), stream_ptr()))
    return float(out.value)


def sam_blend(p_corr, noise, sigma, cond_img, masks, n_masks, gamma=0.5, gain=2.0, cover_max=0.8):
    """The blend loop of _apply_segmentation (ddim.py:818-901) for all frames + the gamma noise blend (:609).
    p_corr, noise [B,C,F,H,W]; sigma float [F]; masks [F,Nmax,H,W] or None; n_masks int [F]."""
    p_corr = _f32c(p_corr, "p_corr"); noise = _f32c(noise, "noise")
    B, Cc, F, H, W = p_corr.shape
    Nmax = 0
    if masks is not None:
        masks = _f32c(masks, "masks"); cond_img = _f32c(cond_img, "cond_img")
        assert masks.shape[0] == F and masks.shape[-2:] == (H, W) and cond_img.shape == (B, Cc, 1, H, W)
        Nmax = masks.shape[1]
    out = torch.empty_like(p_corr)
    sg, nm = host_f32(sigma), host_i32(n_masks)
    with torch.cuda.device(p_corr.device):
        check(lib().moca_sam_blend(ctx(p_corr.device.index), ptr(p_corr), ptr(noise), C.cast(sg, C.c_void_p), ptr(cond_img),
                                   ptr(masks), C.cast(nm, C.c_void_p), Nmax, float(gamma), float(gain), float(cover_max), B, Cc, F,
                                   H, W, ptr(out), stream_ptr()))
    return out


def p_sample_ddim(x, e_c, e_u, noise, cfg_scale, A, A_prev, sigma, sqrt_one_minus_A, scale=None, scale_prev=None):
    """DDIMSampler.p_sample_ddim (ddim.py:339-357), scalar table index; ``scale`` given = use_scale branch."""
    x = _f32c(x, "x"); e_c = _f32c(e_c, "e_c"); noise = _f32c(noise, "noise")
    e_u = _f32c(e_u, "e_u") if e_u is not None else None
    xp = torch.empty_like(x); p0 = torch.empty_like(x)
    use_scale = scale is not None
    with torch.cuda.device(x.device):
        check(lib().moca_p_sample_ddim(ctx(x.device.index), ptr(x), ptr(e_c), ptr(e_u), ptr(noise), float(cfg_scale), float(A),
                                       float(A_prev), float(sigma), float(sqrt_one_minus_A), int(use_scale),
                                       float(scale if use_scale else 1.0), float(scale_prev if use_scale else 1.0), x.numel(),
                                       ptr(xp), ptr(p0), stream_ptr()))
    return xp, p0


def renoise(z, noise, src_frame, alpha):
    """prepare_latents re-noising (scripts/evaluation/funcs.py:55-80): queue[q] = a^.5 z[src[q]] + (1-a)^.5 noise[q]."""
    z = _f32c(z, "z"); noise = _f32c(noise, "noise")
    B, Cc, Fz, H, W = z.shape
    Q = noise.shape[2]
    out = torch.empty_like(noise)
    src_h, alpha_h = host_i32(src_frame), host_f32(alpha)
    with torch.cuda.device(z.device):
        check(lib().moca_renoise(ctx(z.device.index), ptr(z), ptr(noise), C.cast(src_h, C.c_void_p),
                                 C.cast(alpha_h, C.c_void_p), B, Cc, Fz, Q, H, W, ptr(out), stream_ptr()))
    return out


def queue_shift(queue, tail):
    """shift_latents (funcs.py:89-97): drop frame 0, append ``tail`` [B,C,1,H,W]."""
    queue = _f32c(queue, "queue"); tail = _f32c(tail, "tail")
    B, Cc, Q, H, W = queue.shape
    out = torch.empty_like(queue)
    with torch.cuda.device(queue.device):
        check(lib().moca_queue_shift(ctx(queue.device.index), ptr(queue), ptr(tail), B, Cc, Q, H, W, ptr(out), stream_ptr()))
    return out


def window_gather(queue, start, Fw):
    queue = _f32c(queue, "queue")
    B, Cc, Q, H, W = queue.shape
    win = torch.empty((B, Cc, Fw, H, W), device=queue.device, dtype=torch.float32)
    with torch.cuda.device(queue.device):
        check(lib().moca_window_gather(ctx(queue.device.index), ptr(queue), B, Cc, Q, H, W, int(start), int(Fw), ptr(win), stream_ptr()))
    return win


def window_scatter(queue, win, win_off, start, Fw):
    """queue[:, :, start:start+Fw] = win[:, :, win_off:win_off+Fw] in place (funcs.py:351-352)."""
    assert queue.is_cuda and queue.dtype == torch.float32 and queue.is_contiguous()
    win = _f32c(win, "win")
    B, Cc, Q, H, W = queue.shape
    with torch.cuda.device(queue.device):
        check(lib().moca_window_scatter(ctx(queue.device.index), ptr(win), win.shape[2], int(win_off), B, Cc, Q, H, W, int(start),
                                        int(Fw), ptr(queue), stream_ptr()))
    return queue


FILTER_TYPES = {"gaussian": 0, "butterworth": 1, "ideal": 2, "box": 3}


def freeinit_filter(shape, device, filter_type="gaussian", n=4, d_s=0.25, d_t=0.25):
    """get_freq_filter (utils/freeinit_utils.py:51-70) evaluated on the device; returns [T,H,W] fp32."""
    if filter_type not in FILTER_TYPES:
        raise NotImplementedError(filter_type)
    T, H, W = shape[-3], shape[-2], shape[-1]
    dev = torch.device(device)
    if dev.type != "cuda":
        raise _lib.MocaError("freeinit_filter: the filter is built on a CUDA device (no CPU path exists)")
    lpf = torch.empty((T, H, W), device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        check(lib().moca_freeinit_filter(ctx(dev.index), FILTER_TYPES[filter_type], int(n), float(d_s), float(d_t), T, H, W,
                                         ptr(lpf), stream_ptr()))
    return lpf


def freq_mix_3d(x, noise, lpf):
    """freq_mix_3d (utils/freeinit_utils.py:7-47) over the last three dims of x / noise; lpf broadcastable [T,H,W]."""
    x = _f32c(x, "x"); noise = _f32c(noise, "noise")
    T, H, W = x.shape[-3:]
    lpf = _f32c(lpf, "lpf").reshape(-1, T, H, W)[0].contiguous() if lpf.numel() != T * H * W else _f32c(lpf, "lpf").reshape(T, H, W)
    NV = x.numel() // (T * H * W)
    y = torch.empty_like(x)
    work = torch.empty(lib().moca_freeinit_work_bytes(NV, T, H, W), device=x.device, dtype=torch.uint8)
    with torch.cuda.device(x.device):
        check(lib().moca_freeinit_mix(ctx(x.device.index), ptr(x), ptr(noise), ptr(lpf), NV, T, H, W, ptr(y), ptr(work), stream_ptr()))
    return y


def freq_mix_3d_static(x, noise, lpf, out):
    """freq_mix_3d through moca_freeinit_mix_static: no allocation anywhere on the way (``out`` and contiguous fp32 inputs are
    the caller's; the spectrum scratch is the library's built-in 2 MiB array, so NV*T*H*W <= 262144).  No counterpart in the
    reference.  Calls that overlap on the GPU share the scratch and race: keep them on one stream."""
    for t, name in ((x, "x"), (noise, "noise"), (lpf, "lpf"), (out, "out")):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise _lib.MocaError(f"freq_mix_3d_static: {name} must be a contiguous fp32 CUDA tensor (this entry converts nothing)")
    T, H, W = x.shape[-3:]
    if noise.shape != x.shape or out.shape != x.shape or lpf.numel() != T * H * W:
        raise _lib.MocaError("freq_mix_3d_static: x / noise / out must have one shape and lpf T*H*W elements")
    NV = x.numel() // (T * H * W)
    with torch.cuda.device(x.device):
        check(lib().moca_freeinit_mix_static(ctx(x.device.index), ptr(x), ptr(noise), ptr(lpf), NV, T, H, W, ptr(out), stream_ptr()))
    return out


# ---- unit-test primitives -------------------------------------------------------------------------------------
def gemm_op16(a, w, bias=None, out_op16=False):
    M, K = a.shape
    N = w.shape[0]
    out = torch.empty((M, N), device=a.device, dtype=op_dtype() if out_op16 else torch.float32)
    with torch.cuda.device(a.device):
        check(lib().moca_gemm_op16(ctx(a.device.index), ptr(a), ptr(w), ptr(bias), M, N, K, ptr(out), int(out_op16), stream_ptr()))
    return out


def geglu(a_bf16, w, bias):
    M, K = a_bf16.shape
    N = w.shape[0]
    out = torch.empty((M, N // 2), device=a_bf16.device, dtype=op_dtype())
    with torch.cuda.device(a_bf16.device):
        check(lib().moca_geglu_f32(ctx(a_bf16.device.index), ptr(a_bf16), ptr(w), ptr(bias), M, N, K, ptr(out), stream_ptr()))
    return out


def conv3x3_nhwc(x, w, bias, stride=1):
    NF, H, W, Cin = x.shape
    Cout = w.shape[0]
    y = torch.empty((NF, H // stride, W // stride, Cout), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        check(lib().moca_conv3x3_f32(ctx(x.device.index), ptr(x), ptr(w), ptr(bias), NF, H, W, Cin, Cout, stride, ptr(y), stream_ptr()))
    return y


def tconv_tokens(x, w, bias):
    B, T, HW, Cin = x.shape
    Cout = w.shape[0]
    y = torch.empty((B, T, HW, Cout), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        check(lib().moca_tconv_f32(ctx(x.device.index), ptr(x), ptr(w), ptr(bias), B, T, HW, Cin, Cout, ptr(y), stream_ptr()))
    return y


def groupnorm_op16(x, gamma, beta, rows_per_stat, eps, silu):
    rows, Cc = x.shape
    out = torch.empty((rows, Cc), device=x.device, dtype=op_dtype())
    with torch.cuda.device(x.device):
        check(lib().moca_groupnorm_op16(ctx(x.device.index), ptr(x), ptr(gamma), ptr(beta), rows, Cc, rows_per_stat, float(eps),
                                        int(silu), ptr(out), stream_ptr()))
    return out


def layernorm_op16(x, gamma, beta):
    rows, Cc = x.shape
    out = torch.empty((rows, Cc), device=x.device, dtype=op_dtype())
    with torch.cuda.device(x.device):
        check(lib().moca_layernorm_op16(ctx(x.device.index), ptr(x), ptr(gamma), ptr(beta), rows, Cc, ptr(out), stream_ptr()))
    return out


def attention_op16(q, k, v, batch, heads, Lq, Lk, Lk_valid=None, temporal_inner=0):
    """q [batch*Lq, heads*64], k/v [batch*Lk, heads*64] bf16 (row pitch = tensor stride(0))."""
    out = torch.empty((q.shape[0], heads * 64), device=q.device, dtype=op_dtype())
    with torch.cuda.device(q.device):
        check(lib().moca_attention_op16(ctx(q.device.index), ptr(q), ptr(k), ptr(v), batch, heads, Lq, Lk,
                                        Lk if Lk_valid is None else Lk_valid, q.stride(0), k.stride(0), ptr(out), out.stride(0),
                                        int(temporal_inner), stream_ptr()))
    return out


# ---- multi-GPU queue hand-off (include/moca_b200.h: moca_comm_*, moca_queue_handoff) ---------------------------------------
def comm_init(group=None, device=None):
    """Creates the library's NCCL communicator for the ranks of ``group`` (default: the world).  torch.distributed only carries
    the 128-byte NCCL id from rank 0 to the others; the exchange itself (queue_handoff) is issued by the library."""
    import torch.distributed as dist
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    buf = C.create_string_buffer(128)
    if rank == 0:
        check(lib().moca_comm_unique_id(buf))
    box = [bytes(buf.raw)]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    idb = C.create_string_buffer(box[0], 128)
    with torch.cuda.device(dev):
        check(lib().moca_comm_init(ctx(dev.index), idb, world, rank))
    return world


def comm_size(device=None):
    """Ranks of the library communicator of this device's context (0: none)."""
    idx = torch.cuda.current_device() if device is None else torch.device(device).index
    return int(lib().moca_comm_size(ctx(idx)))


def queue_handoff(packed, gathered):
    """All-gather of the ranks' updated half-windows on the current stream (replaces the mp.Queue exchange of funcs_mp.py:197-236).
    packed: fp32 [count...] this rank's contribution; gathered: fp32 [nranks, count...]."""
    assert packed.is_cuda and gathered.is_cuda and packed.dtype == gathered.dtype == torch.float32
    assert packed.is_contiguous() and gathered.is_contiguous() and gathered.numel() % packed.numel() == 0
    with torch.cuda.device(packed.device):
        check(lib().moca_queue_handoff(ctx(packed.device.index), ptr(packed), ptr(gathered), packed.numel(), stream_ptr()))
    return gathered


def comm_destroy(device=None):
    idx = torch.cuda.current_device() if device is None else torch.device(device).index
    check(lib().moca_comm_destroy(ctx(idx)))

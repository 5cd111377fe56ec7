"""Host-side mirror of the FIFO diagonal-denoising driver (scripts/evaluation/funcs.py) for one or several GPUs.

  prepare_latents      funcs.py:21-82     -> moca_renoise
  shift_latents        funcs.py:86-99     -> moca_freeinit_filter / moca_freeinit_mix / moca_queue_shift
  fifo_ddim_sampling   funcs.py:243-371   window loop; windows of one iteration only depend on the iteration-start
                                          queue (funcs.py:306,351-352; scripts/evaluation/funcs_mp.py:207-221), so they
                                          are sharded over the ranks of a torch.distributed group: window r runs on rank
                                          r mod G, the updated half-windows are exchanged with ONE all_gather per
                                          iteration (NCCL over NVLink), every rank then applies the same shift.
RNG parity: the noise stream is the reference's (window 2n-1 .. 0, one draw per frame, then the shift's tail draw) from
an identically seeded generator on every rank; a rank draws only the windows it owns and advances the generator's
Philox offset past the others (``_NoiseStream``), so 1-GPU and G-GPU runs produce the same queue without 8x the launches.  Popped latent frames are returned (or handed to ``on_frame``: driver.run_prompts decodes them with the library's VAE
decoder, vae.py, and writes the images on a worker thread).
"""
import numpy as np
import torch

from . import ops
from .schedule import fifo_queue_tables


def prepare_latents(args, input_path, sampler, model=None, data=None, initial_latents=None, noise=None, generator=None):
    """funcs.py:21-82.  ``initial_latents`` replaces the reference's torch.load(f"{input_path}/{S}.pt") when given;
    ``noise`` [B,C,Q,H,W] replaces the per-slot randn_like draws (same order: look-ahead copies first)."""
    if data is not None:
        # DAVIS branch (funcs.py:38-47): the queue starts from the encoded video frames instead of the cached base sample
        if model is None or not hasattr(model, "encode_first_stage_2DAE"):
            raise ValueError("prepare_latents: data=(frames, masks) needs a model with encode_first_stage_2DAE (the VAE encoder)")
        initial_latents = model.encode_first_stage_2DAE(_rgb(data[0]).to("cuda"))
    if initial_latents is None:
        initial_latents = torch.load(f"{input_path}/{args.num_inference_steps}.pt")
    z = initial_latents.to("cuda", dtype=torch.float32).contiguous()
    B, C, Fz, H, W = z.shape
    S = args.num_inference_steps
    look = args.video_length // 2 if args.lookahead_denoising else 0
    Q = look + S
    if noise is None:
        noise = torch.cat([torch.randn((B, C, 1, H, W), device=z.device, generator=generator) for _ in range(Q)], dim=2)
    src = [0] * look + [max(0, i - (S - Fz)) for i in range(S)]
    alpha = [float(sampler.ddim_alphas[0])] * look + [float(sampler.ddim_alphas[i]) for i in range(S)]
    return ops.renoise(z, noise, src, alpha)


def _rgb(frames):
    """funcs.py:43-44 / :103-104: RGBA frames lose their alpha channel."""
    return frames[:, :3] if frames.shape[1] == 4 else frames


_filter_cache = {}


def shift_latents(latents, davis_data=None, model=None, new_noise=None, generator=None):
    """funcs.py:86-99 (non-DAVIS branch): anchor = frame 0; queue shifts left by one; the new tail frame is
    freq_mix_3d(anchor, N(0,1), gaussian LPF(d_s=d_t=0.25)) built for a T=1 volume."""
    B, C, Q, H, W = latents.shape
    if davis_data is not None:
        # DAVIS branch (funcs.py:100-116): the anchor is the encoded LAST video frame, re-encoded every call like the reference
        # (its posterior sample draws fresh host noise each time); the mask queue shifts with the latents and returns with them
        if model is None or not hasattr(model, "encode_first_stage_2DAE"):
            raise ValueError("shift_latents: davis_data needs a model with encode_first_stage_2DAE (the VAE encoder)")
        frames, masks = davis_data
        anchor = model.encode_first_stage_2DAE(_rgb(frames[:, :, -1:]).to(latents.device)).to(torch.float32).contiguous()
    else:
        anchor = ops.window_gather(latents, 0, 1)
    if new_noise is None:
        new_noise = torch.randn((B, C, H, W), device=latents.device, generator=generator).unsqueeze(2)
    key = (H, W, latents.device.index)
    if key not in _filter_cache:      # the reference rebuilds the filter on the CPU every call (freeinit_utils.py:73-91)
        _filter_cache[key] = ops.freeinit_filter((1, H, W), latents.device, "gaussian", 1, 0.25, 0.25)
    tail = ops.freq_mix_3d(anchor, new_noise.reshape(B, C, 1, H, W), _filter_cache[key])
    out = ops.queue_shift(latents, tail)
    if davis_data is not None:
        if masks is not None:
            masks = torch.cat([masks[:, :, 1:], masks[:, :, -1:]], dim=2)     # funcs.py:114-115 (the tail keeps its own value)
        return out, (frames, masks)
    return out


class _NoiseStream:
    """The reference's per-frame N(0,1) draws (ddim.py:561 through lvdm/common.py:31-34; funcs.py:93) as a stream that can
    be skipped: ``block(n, keep)`` returns n frames drawn one randn call each, or advances the generator past n such calls.
    CUDA generators are advanced through their Philox offset (measured from one real call); CPU generators (the gloo
    tests) draw and discard."""

    def __init__(self, generator, device, frame_shape):
        self.dev = torch.device(device)
        self.shape = tuple(frame_shape)
        self.gen = generator
        if generator is None and self.dev.type == "cuda":
            self.gen = torch.cuda.default_generators[self.dev.index if self.dev.index is not None else torch.cuda.current_device()]
        self.delta = None

    def _one(self):
        if self.dev.type == "cuda" and self.delta is None:
            o0 = self.gen.get_offset()
            t = torch.randn(self.shape, device=self.dev, generator=self.gen)
            self.delta = self.gen.get_offset() - o0
            return t
        return torch.randn(self.shape, device=self.dev, generator=self.gen)

    def block(self, n, keep=True):
        if keep:
            return torch.cat([self._one() for _ in range(n)], dim=2)
        if self.dev.type != "cuda":
            for _ in range(n):
                self._one()
            return None
        if self.delta is None:
            self._one()
            n -= 1
        self.gen.set_offset(self.gen.get_offset() + n * self.delta)
        return None


_xchg_cache = {}


def _xchg_buffers(key, per, world, shape, dev):
    """Hand-off staging, allocated once per geometry (no allocator traffic between the graph replays of an iteration)."""
    if key not in _xchg_cache:
        _xchg_cache[key] = (torch.zeros((per,) + shape, device=dev, dtype=torch.float32),
                            torch.empty((world, per) + shape, device=dev, dtype=torch.float32))
    return _xchg_cache[key]


def window_plan(args):
    """[(start, mid, end)] for rank = 2n-1 .. 0 in the reference's (reversed) order, funcs.py:306-310."""
    f = args.video_length
    n = 2 * args.num_partitions if args.lookahead_denoising else args.num_partitions
    out = []
    for rank in reversed(range(n)):
        start = rank * (f // 2) if args.lookahead_denoising else rank * f
        out.append((rank, start, start + f // 2, start + f))
    return out


def fifo_ddim_sampling(args, model, conditioning, noise_shape, ddim_sampler, cfg_scale=1.0, output_dir=None, latents_dir=None,
                       save_frames=False, conditioned_image=None, targets=None, gamma=0.5, use_self_attention=False,
                       davis_data=None, anchor_frame=None, unconditional_conditioning=None, masks=None, latents=None,
                       generator=None, group=None, on_frame=None, num_iterations=None, queue_out=None, shard=True,
                       **kwargs):
    """funcs.py:243-371.  Returns the list of popped latent frames [B,C,1,H,W] (one per iteration).

    ``unconditional_conditioning``: the reference encodes "" with CLIP here (funcs.py:267-276); CLIP is out of scope, so
    the embedding is an input (or ``model.get_learned_conditioning`` is used when the model has it).
    ``masks`` [B,1,Q,H,W]: per-queue-slot masks (the reference's DAVIS masks); ``davis_data`` = (frames [B,3|4,T,H*8,W*8], masks): the
    reference's DAVIS mode — the queue starts from the encoded frames and the shift anchors on the encoded last frame (VAE encoder).  ``group``: torch.distributed group to
    shard windows over (None = the default group if initialised, else single GPU).  ``queue_out``: dict that receives the
    final queue (``latents``, ``masks``) so a caller can continue the run.  ``shard=False`` runs every window on the calling
    rank even when a process group exists (the 1-GPU replay the multi-GPU equality checks compare with)."""
    import torch.distributed as dist
    kwargs.update({"clean_cond": True})
    cond = conditioning
    uc = None
    if cfg_scale != 1.0:
        uc = unconditional_conditioning
        if uc is None:
            uc_emb = model.get_learned_conditioning(noise_shape[0] * [""])
            uc = dict(cond)
            uc["c_crossattn"] = [uc_emb]
    if latents is None:
        latents = prepare_latents(args, latents_dir, ddim_sampler, model, data=davis_data, generator=generator)
    if davis_data is not None and masks is None and davis_data[1] is not None:
        masks = davis_data[1].to(latents.device, dtype=torch.float32).contiguous()       # funcs.py:297-301
        davis_data = (davis_data[0], masks)
    B, C, Q, H, W = latents.shape
    f = args.video_length
    half = f // 2
    ts_q, idx_q = fifo_queue_tables(args.num_inference_steps, f, args.lookahead_denoising)
    wins = window_plan(args)
    world = dist.get_world_size(group) if shard and (group is not None or (dist.is_available() and dist.is_initialized())) else 1
    rank = dist.get_rank(group) if world > 1 else 0
    n_iter = num_iterations if num_iterations is not None else args.new_video_length + args.num_inference_steps - f
    popped = []
    dev = latents.device
    stream = _NoiseStream(generator, dev, (B, C, 1, H, W))
    keep = half if args.lookahead_denoising else f
    per = (len(wins) + world - 1) // world

    def onestep(start, end, nz):
        x_in = ops.window_gather(latents, start, f)
        m_in = ops.window_gather(masks, start, f) if masks is not None else None
        return ddim_sampler.fifo_onestep(cond=cond, shape=noise_shape, latents=x_in, timesteps=ts_q[start:end],
                                         indices=idx_q[start:end], unconditional_guidance_scale=cfg_scale,
                                         unconditional_conditioning=uc, cond_image=conditioned_image, target=targets,
                                         gamma=gamma, use_self_attention=use_self_attention, davis_masks=m_in, noise=nz,
                                         **kwargs)[0]

    for it in range(n_iter):
        if world == 1:
            for (w, start, mid, end) in wins:
                x_prev = onestep(start, end, stream.block(f))
                if args.lookahead_denoising:
                    ops.window_scatter(latents, x_prev, half, mid, half)
                else:
                    ops.window_scatter(latents, x_prev, 0, start, f)
        else:
            # snapshot semantics (funcs_mp.py:207): all windows read the iteration-start queue, so the owned windows run
            # first and the queue is only written after the exchange
            packed, gathered = _xchg_buffers((world, per, B, C, keep, H, W, dev.index), per, world, (B, C, keep, H, W), dev)
            j = 0
            for (w, start, mid, end) in wins:
                nz = stream.block(f, keep=(w % world == rank))
                if nz is None:
                    continue
                x_prev = onestep(start, end, nz)
                ops.window_scatter(packed[j], x_prev, f - keep, 0, keep)
                j += 1
            # the queue hand-off: issued by the library (moca_queue_handoff, NCCL on the current stream) once
            # ops.comm_init(group) has created its communicator; torch.distributed carries it otherwise (gloo / CPU tests)
            if getattr(ops, "comm_size", None) is not None and latents.is_cuda and ops.comm_size(dev) == world:
                ops.queue_handoff(packed, gathered)
            else:
                dist.all_gather_into_tensor(gathered.view(-1), packed.view(-1), group=group)
            for (w, start, mid, end) in wins:
                owner = w % world
                slot = [x[0] for x in wins if x[0] % world == owner].index(w)       # same order as on the owner
                ops.window_scatter(latents, gathered[owner, slot], 0, mid if args.lookahead_denoising else start, keep)
        first = half if args.lookahead_denoising else 0
        frame = ops.window_gather(latents, first, 1)
        if on_frame is not None:
            on_frame(it, frame)
        popped.append(frame)
        # funcs.py:93 draws randn_like(latents[:, :, -1]) = [B,C,H,W]; same element count and Philox consumption as one frame
        if davis_data is not None and masks is not None:
            latents, davis_data = shift_latents(latents, davis_data, model, new_noise=stream.block(1))       # funcs.py:368-369
            masks = davis_data[1]
        else:
            latents = shift_latents(latents, new_noise=stream.block(1))
            if masks is not None:
                masks = ops.queue_shift(masks, ops.window_gather(masks, Q - 1, 1))
    if queue_out is not None:
        queue_out["latents"], queue_out["masks"] = latents, masks
    return popped

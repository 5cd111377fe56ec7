"""Driver / IO around the hot path: what ``videocrafter_main.py:176-232`` does around the sampler, kept off
the critical path of the FIFO loop.

* ``load_prompts``        prompt CSV (scripts/evaluation/funcs.py:506-535): same columns, same ``prompt_index`` semantics.
* ``shard``               the reference's ``indices[args.rank::args.num_processes]`` (videocrafter_main.py:181), with rank / world
                          taken from the torchrun environment when not given.
* ``set_directory``       output / latent-cache directory naming (videocrafter_main.py:25-62).
* ``latent_cache_*``      the ``{latents_dir}/0.pt`` / ``{S}.pt`` cache that decides whether the base DDIM run is needed
                          (videocrafter_main.py:201, lvdm/models/samplers/ddim.py:233-250); sampler.DDIMSampler.ddim_sampling writes
                          the same two files.
* ``FrameWriter``         frame -> uint8 image (funcs.py:626-637 ``tensor2image``) and PNG / GIF writing on a worker thread, so the
                          device loop never waits for the file system (the reference saves inside the loop, funcs.py:359-365).
* ``run_prompts``         the per-prompt loop of ``main`` on this package's sampler + FIFO queue, for a model object that offers the
                          reference's ``get_learned_conditioning`` / ``decode_first_stage_2DAE`` (the decoder is the library's,
                          vae.py / row N1, after ``patch_reference()`` or as ``LatentDiffusionB200(first_stage=...)``; the text
                          tower, row N4, is the reference's own module).
* ``main``                ``python -m moca_video_b200.driver`` / torchrun launcher: patches the reference tree and calls its ``main``
                          unchanged, one process per GPU; ``--parallel windows`` shards the FIFO windows over the ranks (NCCL
                          hand-off inside the library), ``--parallel prompts`` shards the prompt list like the reference.

Nothing here computes on the host what the library computes on the device; image conversion of popped frames is host work in the
reference too.
"""
import csv
import os
import queue
import threading

import numpy as np
import torch

from ._lib import MocaError


def load_prompts(prompt_file, prompt_index=None):
    """scripts/evaluation/funcs.py:506-535."""
    rows = []
    with open(prompt_file, "r") as f:
        for row in csv.DictReader(f):
            rows.append({"prompt": row["prompt"].strip(),
                         "conditioned_object": row["conditioned_object"].strip(),
                         "conditioned_image_path": row["conditioned_image_path"].strip(),
                         "conditioned_prompt": row["conditioned_prompt"].strip() + ".",
                         "gamma": float(row["gamma"].strip())})
    if prompt_index is not None:
        if prompt_index >= len(rows):
            raise ValueError(f"Prompt index {prompt_index} exceeds number of available prompts")
        return [rows[prompt_index]]
    return rows


def shard(n_items, rank=None, world=None):
    """``indices[rank::world]`` (videocrafter_main.py:179-181); defaults from RANK / WORLD_SIZE (torchrun)."""
    rank = int(os.environ.get("RANK", "0")) if rank is None else int(rank)
    world = int(os.environ.get("WORLD_SIZE", "1")) if world is None else int(world)
    if not 0 <= rank < world:
        raise ValueError(f"shard: rank {rank} outside 0..{world - 1}")
    return list(range(n_items))[rank::world]


def set_directory(args, prompt, conditioned_image_path=None, create=True):
    """videocrafter_main.py:25-62: (output_dir, latents_dir) for a prompt."""
    p = prompt[:100]
    if args.output_dir is None:
        mode = "self_attention" if getattr(args, "use_self_attention", False) else "sam2"
        output_dir = f"results/videocraft_v2_fifo/random_noise/{mode}/{p}"
        if args.eta != 1.0:
            output_dir += f"/eta{args.eta}"
        if args.new_video_length != 100:
            output_dir += f"/{args.new_video_length}frames"
        if not args.lookahead_denoising:
            output_dir = output_dir.replace(p, f"{p}/no_lookahead_denoising")
        if args.num_partitions != 4:
            output_dir = output_dir.replace(p, f"{p}/n={args.num_partitions}")
        if args.video_length != 16:
            output_dir = output_dir.replace(p, f"{p}/f={args.video_length}")
    else:
        output_dir = args.output_dir
    if getattr(args, "use_davis", False):
        latents_dir = f"visualizations/davis_data/{args.video_name}"
    else:
        latents_dir = f"results/videocraft_v2_fifo/latents/{args.num_inference_steps}steps/{p}/eta{args.eta}"
    if create:
        os.makedirs(output_dir, exist_ok=True)
        os.makedirs(latents_dir, exist_ok=True)
    if conditioned_image_path is not None:
        name = conditioned_image_path.split("/")[-1]
        output_dir = (args.output_dir + "/" + name) if getattr(args, "use_davis", False) else (output_dir + "/" + name.split(".")[0])
        if create:
            os.makedirs(output_dir, exist_ok=True)
    return output_dir, latents_dir


def latent_cache_paths(latents_dir, num_inference_steps):
    return os.path.join(latents_dir, "0.pt"), os.path.join(latents_dir, f"{num_inference_steps}.pt")


def latent_cache_ready(latents_dir, num_inference_steps):
    """``not is_run_base`` of videocrafter_main.py:201."""
    a, b = latent_cache_paths(latents_dir, num_inference_steps)
    return os.path.exists(a) and os.path.exists(b)


def tensor_to_u8(frame):
    """funcs.py:626-637 ``tensor2image`` up to the PIL object: decoded frame [..., 3, H, W] in [-1, 1] -> uint8 [H, W, 3]
    (clamp, (x + 1) / 2, * 255, truncating cast — the reference's own arithmetic)."""
    img = torch.squeeze(frame).detach().float().cpu()
    img = torch.clamp(img, -1.0, 1.0)
    img = (img + 1.0) / 2.0
    return (img * 255).to(torch.uint8).permute(1, 2, 0).contiguous().numpy()


class FrameWriter:
    """Image conversion and file writes of popped frames on a worker thread: ``put`` only enqueues (a device tensor is copied to
    the host on the worker, after the producing stream's event), so the FIFO loop is never blocked by PNG encoding.

    ``frames`` keeps the uint8 images in submission order (what videocrafter_main.py:226-229 hands to the video writer)."""

    def __init__(self, out_dir=None, save_frames=True, prefix=""):
        self.out_dir, self.save_frames, self.prefix = out_dir, save_frames and out_dir is not None, prefix
        if self.save_frames:
            os.makedirs(out_dir, exist_ok=True)
        self.frames = []
        self._q = queue.Queue()
        self._err = None
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def _run(self):
        while True:
            item = self._q.get()
            if item is None:
                return
            i, frame, ev = item
            try:
                if ev is not None:
                    ev.synchronize()
                img = tensor_to_u8(frame)
                while len(self.frames) <= i:
                    self.frames.append(None)
                self.frames[i] = img
                if self.save_frames:
                    from PIL import Image
                    Image.fromarray(img).save(os.path.join(self.out_dir, f"{self.prefix}{i}.png"))
            except Exception as e:  # noqa: BLE001
                self._err = e

    def put(self, i, frame):
        ev = None
        if isinstance(frame, torch.Tensor) and frame.is_cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(frame.device))
        self._q.put((int(i), frame, ev))

    def close(self):
        self._q.put(None)
        self._t.join()
        if self._err is not None:
            raise self._err
        return self.frames


def write_video(path, frames, fps=10, use_mp4=True):
    """videocrafter_main.py:226-229: mp4 (or gif) of uint8 frames through imageio; a GIF through PIL when imageio is absent."""
    try:
        import imageio
        if use_mp4:
            imageio.mimsave(path + ".mp4", frames, fps=fps)
            return path + ".mp4"
        imageio.mimsave(path + ".gif", frames, duration=int(1000 / fps))
        return path + ".gif"
    except ImportError:
        from PIL import Image
        ims = [Image.fromarray(f) for f in frames]
        ims[0].save(path + ".gif", save_all=True, append_images=ims[1:], duration=int(1000 / fps), loop=0)
        return path + ".gif"


def run_prompts(args, model, sampler_cls=None, cond_image=None, decode=None, parallel="windows", group=None, log=print):
    """The prompt loop of videocrafter_main.py:176-232 on this package's sampler and FIFO queue.

    ``model``: offers ``get_learned_conditioning(list[str])`` (text tower) and, unless ``decode`` is given,
    ``decode_first_stage_2DAE(z)`` (VAE, N1) — the reference's LatentDiffusion after ``patch_reference()`` (its decode is then
    the library's), or ``LatentDiffusionB200(first_stage=AutoencoderKLDecoder(...))`` with a text encoder attached.  ``parallel``: "windows" — every rank works on every prompt and
    the FIFO windows of an iteration are sharded over the ranks (fifo.fifo_ddim_sampling); "prompts" — the reference's
    ``indices[rank::num_processes]`` split, each rank runs whole prompts alone.  Returns {prompt index: video path} of the
    prompts this rank wrote."""
    from . import fifo
    from .sampler import DDIMSampler
    import torch.distributed as dist
    sampler_cls = sampler_cls or DDIMSampler
    decode = decode or model.decode_first_stage_2DAE
    prompts = load_prompts(args.prompt_file, getattr(args, "prompt_index", None))
    distributed = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if distributed else int(getattr(args, "rank", 0))
    world = dist.get_world_size(group) if distributed else int(getattr(args, "num_processes", 1))
    todo = shard(len(prompts), rank, world) if parallel == "prompts" else list(range(len(prompts)))
    channels = 4
    h, w, f = args.height // 8, args.width // 8, args.video_length
    written = {}
    for idx in todo:
        d = prompts[idx]
        output_dir, latents_dir = set_directory(args, d["prompt"], d["conditioned_image_path"], create=(rank == 0 or parallel == "prompts"))
        if distributed and parallel == "windows":
            dist.barrier(group)                                      # directories exist before anybody reads the cache
        noise_shape = [1, channels, f, h, w]
        fps = torch.tensor([args.fps], device=model.device).long()
        cond = {"c_crossattn": [model.get_learned_conditioning([d["prompt"]])], "fps": fps}
        sampler = sampler_cls(model)
        if latent_cache_ready(latents_dir, args.num_inference_steps):
            sampler.make_schedule(ddim_num_steps=args.num_inference_steps, ddim_eta=args.eta, verbose=False)
        else:
            # base_ddim_sampling (funcs.py:177-236): one plain DDIM run that fills 0.pt / {S}.pt; rank 0 writes the cache
            uc = {"c_crossattn": [model.get_learned_conditioning([""])], "fps": fps}
            write = latents_dir if (rank == 0 or parallel == "prompts") else None
            sampler.sample(S=args.num_inference_steps, batch_size=1, shape=noise_shape[1:], conditioning=cond, eta=args.eta,
                           unconditional_guidance_scale=args.unconditional_guidance_scale, unconditional_conditioning=uc,
                           latents_dir=write, generator=_seeded(model.device, args.seed))
            if distributed and parallel == "windows":
                dist.barrier(group)
        if d["conditioned_prompt"]:
            cond["c_crossattn"].append(model.get_learned_conditioning([d["conditioned_prompt"]]))
        keep_frames = rank == 0 or parallel == "prompts"
        writer = FrameWriter(os.path.join(output_dir, "fifo") if keep_frames else None, save_frames=bool(args.save_frames) and keep_frames)

        # the popped frame is decoded on a side stream: the next iteration's windows do not wait for the VAE (2.4 ms of a 40 ms
        # iteration at 8 GPUs); the frame tensor is a copy the queue never writes again
        side = torch.cuda.Stream(device=model.device) if keep_frames and torch.device(model.device).type == "cuda" else None

        def on_frame(i, z):
            if not keep_frames:
                return
            if side is None:
                writer.put(i, decode(z))
                return
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(z.device))
            with torch.cuda.stream(side):
                side.wait_event(ready)
                z.record_stream(side)
                writer.put(i, decode(z))                             # FrameWriter records its event on the side stream
        fifo.fifo_ddim_sampling(args, model, cond, noise_shape, sampler, cfg_scale=args.unconditional_guidance_scale,
                                output_dir=output_dir, latents_dir=latents_dir, conditioned_image=cond_image,
                                targets=d["conditioned_object"] + ".", gamma=d["gamma"],
                                use_self_attention=getattr(args, "use_self_attention", False), on_frame=on_frame,
                                shard=(parallel == "windows"), group=group, generator=_seeded(model.device, args.seed + 1))
        frames = writer.close()
        if keep_frames and frames:
            out = output_dir + "/fifo" if args.output_dir is None else output_dir + f"/{d['prompt'][:100]}"
            written[idx] = write_video(out, frames[-args.new_video_length // 2:], fps=args.output_fps, use_mp4=args.use_mp4)
            log(f"[rank {rank}] prompt {idx}: {written[idx]}")
    return written


def _seeded(device, seed):
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    return g


def main(argv=None):
    """Launcher: ``python -m moca_video_b200.driver [--reference PATH] [--parallel windows|prompts] <videocrafter_main.py flags>``
    (under ``torchrun --nproc-per-node N`` for N GPUs).  Installs the library behind the reference's own classes
    (``patch_reference``), creates the hand-off communicator when there are several ranks, and runs the reference's ``main`` — its
    CLI, checkpoint loading, CLIP and VAE stay the reference's own."""
    import argparse
    import sys
    ap = argparse.ArgumentParser(add_help=False)
    ap.add_argument("--reference", default=os.environ.get("MOCA_REFERENCE", "."))
    ap.add_argument("--parallel", default="windows", choices=["windows", "prompts"])
    own, rest = ap.parse_known_args(argv)
    if not os.path.isdir(os.path.join(own.reference, "lvdm")):
        raise MocaError(f"driver: '{own.reference}' is not a MoCA-Video tree (pass --reference or set MOCA_REFERENCE)")
    sys.path.insert(0, os.path.abspath(own.reference))
    import moca_video_b200 as mb
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    if not torch.cuda.is_available():
        raise MocaError("driver: no CUDA device (the library has no CPU path)")
    torch.cuda.set_device(local)
    mb.patch_reference()
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if own.parallel == "windows":
            from . import ops
            ops.comm_init()                       # FIFO windows of every prompt sharded over the ranks
        else:
            rest += ["--rank", str(rank), "--num_processes", str(world)]      # the reference's own prompt split
    # the reference keeps its argument parser under ``if __name__ == "__main__"``: run the script itself, with the patched modules
    # already in sys.modules
    import runpy
    sys.argv = ["videocrafter_main.py"] + rest
    return runpy.run_path(os.path.join(os.path.abspath(own.reference), "videocrafter_main.py"), run_name="__main__")


if __name__ == "__main__":
    main()

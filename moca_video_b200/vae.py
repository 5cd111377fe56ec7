"""Host-side mirror of the reference first-stage decode path.

``AutoencoderKLDecoder`` keeps the reference's constructor vocabulary (``ddconfig`` of the yaml's ``first_stage_config``,
configs/inference_t2v_512_v2.0.yaml:52-72), takes the reference ``AutoencoderKL.state_dict()`` unchanged (encoder / quant_conv /
loss keys are ignored: they are not on the decode path) and runs ``decode`` (lvdm/models/autoencoder.py:104-107: post_quant_conv +
``Decoder.forward``, lvdm/modules/networks/ae_modules.py:536-578) inside libmoca_b200.so.  ``decode_first_stage_2DAE`` is the
per-frame loop of ``LatentDiffusion.decode_first_stage_2DAE`` (lvdm/models/ddpm3d.py:555-562) with the frames of a call
decoded ``frames_per_call`` at a time (frames are independent in a 2-D autoencoder, so the grouping does not change a value).
"""
import ctypes as C

import torch

from . import _lib
from ._lib import check, ctx, host_i32, lib, ptr, stream_ptr

# ddconfig options of Decoder.__init__ (ae_modules.py:467-470) that select code the library does not implement
_UNSUPPORTED = dict(give_pre_end=False, tanh_out=False, use_linear_attn=False, attn_type="vanilla", resamp_with_conv=True)


def decoder_param_shapes(ch, ch_mult, num_res_blocks, z_channels, out_ch, embed_dim=None):
    """{state_dict key: shape} of post_quant_conv + Decoder, in the reference's registration order (ae_modules.py:488-534)."""
    embed_dim = z_channels if embed_dim is None else embed_dim
    shapes = {"post_quant_conv.weight": (z_channels, embed_dim, 1, 1), "post_quant_conv.bias": (z_channels,)}

    def conv(name, co, ci, k):
        shapes[name + ".weight"] = (co, ci, k, k)
        shapes[name + ".bias"] = (co,)

    def norm(name, c):
        shapes[name + ".weight"] = (c,)
        shapes[name + ".bias"] = (c,)

    def res(name, ci, co):
        norm(name + ".norm1", ci)
        conv(name + ".conv1", co, ci, 3)
        norm(name + ".norm2", co)
        conv(name + ".conv2", co, co, 3)
        if ci != co:
            conv(name + ".nin_shortcut", co, ci, 1)

    L = len(ch_mult)
    block_in = ch * ch_mult[L - 1]
    conv("decoder.conv_in", block_in, z_channels, 3)
    res("decoder.mid.block_1", block_in, block_in)
    norm("decoder.mid.attn_1.norm", block_in)
    for n in ("q", "k", "v", "proj_out"):
        conv(f"decoder.mid.attn_1.{n}", block_in, block_in, 1)
    res("decoder.mid.block_2", block_in, block_in)
    for lvl in reversed(range(L)):
        block_out = ch * ch_mult[lvl]
        for j in range(num_res_blocks + 1):
            res(f"decoder.up.{lvl}.block.{j}", block_in, block_out)
            block_in = block_out
        if lvl != 0:
            conv(f"decoder.up.{lvl}.upsample.conv", block_in, block_in, 3)
    norm("decoder.norm_out", block_in)
    conv("decoder.conv_out", out_ch, block_in, 3)
    return shapes


class AutoencoderKLDecoder:
    def __init__(self, ddconfig=None, embed_dim=4, scale_factor=0.18215, device=None, frames_per_call=2, _listing_only=False,
                 **ignored):
        dd = dict(ddconfig or {})
        for k, want in _UNSUPPORTED.items():
            if k in dd and dd[k] != want:
                raise _lib.MocaError(f"AutoencoderKLDecoder: ddconfig option {k}={dd[k]!r} is not supported (needs {want!r})")
        if dd.get("attn_resolutions"):
            raise _lib.MocaError("AutoencoderKLDecoder: attention inside the up levels (attn_resolutions) is not supported")
        self.ch = int(dd.get("ch", 128))
        self.ch_mult = tuple(int(v) for v in dd.get("ch_mult", (1, 2, 4, 4)))
        self.num_res_blocks = int(dd.get("num_res_blocks", 2))
        self.z_channels = int(dd.get("z_channels", 4))
        self.out_ch = int(dd.get("out_ch", 3))
        if int(embed_dim) != self.z_channels:
            raise _lib.MocaError(f"AutoencoderKLDecoder: embed_dim {embed_dim} != z_channels {self.z_channels} is not supported")
        self.scale_factor = float(scale_factor)
        self.frames_per_call = int(frames_per_call)
        self.shapes = decoder_param_shapes(self.ch, self.ch_mult, self.num_res_blocks, self.z_channels, self.out_ch)
        self.upscale = 2 ** (len(self.ch_mult) - 1)
        self._h = C.c_void_p()
        self.device = None
        if _listing_only:
            c = None
        else:
            self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
            c = ctx(self.device.index)
        mult = host_i32(self.ch_mult)
        check(lib().moca_vae_create(c, self.ch, C.cast(mult, C.c_void_p), len(self.ch_mult), self.num_res_blocks, self.z_channels,
                                    self.out_ch, C.byref(self._h)))
        self._loaded = False

    def expected_params(self):
        """{state_dict key: numel} as the library's loader expects them."""
        n = lib().moca_unet_param_count(self._h)
        out = {}
        buf = C.create_string_buffer(512)
        numel = C.c_int64()
        for i in range(n):
            check(lib().moca_unet_param_info(self._h, i, buf, 512, C.byref(numel)))
            out[buf.value.decode()] = int(numel.value)
        return out

    def load_state_dict(self, sd, strict=True):
        """Takes ``AutoencoderKL.state_dict()`` (or ``first_stage_model.*``-stripped checkpoint keys).  Keys outside the decode
        path (encoder.*, quant_conv.*, loss.*) are skipped; a missing decode-path key raises."""
        missing = [k for k in self.shapes if k not in sd]
        if missing and strict:
            raise _lib.MocaError(f"load_state_dict: missing {missing[:4]} ({len(missing)} keys)")
        with torch.cuda.device(self.device):
            names, tensors = [], []
            for k, shp in self.shapes.items():
                if k not in sd:
                    continue
                t = sd[k]
                if tuple(t.shape) != tuple(shp):
                    raise _lib.MocaError(f"load_state_dict: {k} has shape {tuple(t.shape)}, expected {tuple(shp)}")
                names.append(k.encode())
                tensors.append(t.detach().to(device=self.device, dtype=torch.float32).contiguous())
            n = len(names)
            c_names = (C.c_char_p * n)(*names)
            c_ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in tensors])
            c_numel = (C.c_int64 * n)(*[t.numel() for t in tensors])
            check(lib().moca_unet_load_many(self._h, n, c_names, c_ptrs, c_numel, stream_ptr()))
            del tensors
            check(lib().moca_unet_finalize(self._h))
        self._loaded = True
        return self

    def eval(self):
        return self

    # ---- decode ------------------------------------------------------------------------------------------------
    def _decode(self, z, scale):
        if not self._loaded:
            raise _lib.MocaError("AutoencoderKLDecoder.decode: weights not loaded")
        if z.dim() != 4 or z.shape[1] != self.z_channels:
            raise _lib.MocaError(f"AutoencoderKLDecoder.decode: z must be [n, {self.z_channels}, h, w], got {tuple(z.shape)}")
        n, _, h, w = z.shape
        z = z.to(device=self.device, dtype=torch.float32).contiguous()
        out = torch.empty((n, self.out_ch, h * self.upscale, w * self.upscale), device=self.device, dtype=torch.float32)
        step = max(1, min(self.frames_per_call, 16))
        with torch.cuda.device(self.device):
            for i in range(0, n, step):
                m = min(step, n - i)
                check(lib().moca_vae_decode(self._h, ptr(z[i:i + m]), m, h, w, scale, ptr(out[i:i + m]), stream_ptr()))
        return out

    @torch.no_grad()
    def decode(self, z, **kwargs):
        """AutoencoderKL.decode (autoencoder.py:104-107): z [n, zc, h, w] (already divided by scale_factor) -> [n, 3, 8h, 8w]."""
        return self._decode(z, 1.0)

    @torch.no_grad()
    def decode_first_stage_2DAE(self, z, **kwargs):
        """LatentDiffusion.decode_first_stage_2DAE (ddpm3d.py:555-562): z [b, zc, t, h, w] sampler latents -> [b, 3, t, 8h, 8w];
        the 1 / scale_factor is applied while the latents are staged."""
        b, c, t, h, w = z.shape
        frames = z.permute(0, 2, 1, 3, 4).reshape(b * t, c, h, w)
        img = self._decode(frames, self.scale_factor)
        return img.reshape(b, t, self.out_ch, img.shape[-2], img.shape[-1]).permute(0, 2, 1, 3, 4).contiguous()

    # ---- accounting ----------------------------------------------------------------------------------------------
    def launches(self):
        return int(lib().moca_unet_launches(self._h))

    def flops(self):
        return float(lib().moca_unet_flops(self._h))

    def workspace_bytes(self):
        return int(lib().moca_unet_workspace_bytes(self._h))

    def profile(self, iters=5):
        need = lib().moca_unet_profile(self._h, 1, None, 0, stream_ptr())
        buf = C.create_string_buffer(int(need) + 4096)
        lib().moca_unet_profile(self._h, int(iters), buf, len(buf), stream_ptr())
        rows = []
        for line in buf.value.decode().splitlines():
            i, us, gf, label = line.split(",", 3)
            rows.append((float(us), float(gf), label))
        return rows

    def close(self):
        if self._h:
            lib().moca_unet_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001  (interpreter shutdown)
            pass


def attach_vae(first_stage_model, scale_factor=0.18215, device=None, frames_per_call=2):
    """Builds (once) the library decoder of a reference ``AutoencoderKL`` from its own modules and ``state_dict``."""
    eng = getattr(first_stage_model, "_moca_vae", None)
    if eng is None:
        d = first_stage_model.decoder
        if getattr(d, "give_pre_end", False) or getattr(d, "tanh_out", False):
            raise _lib.MocaError("attach_vae: give_pre_end / tanh_out decoders are not supported")
        if any(len(u.attn) for u in d.up):
            raise _lib.MocaError("attach_vae: attention inside the up levels is not supported")
        ch = int(d.ch)
        mult = tuple(int(u.block[0].conv1.out_channels) // ch for u in d.up)
        dd = dict(ch=ch, ch_mult=mult, num_res_blocks=int(d.num_res_blocks), z_channels=int(d.conv_in.in_channels),
                  out_ch=int(d.conv_out.out_channels))
        dev = device if device is not None else next(first_stage_model.parameters()).device
        eng = AutoencoderKLDecoder(dd, embed_dim=int(first_stage_model.post_quant_conv.in_channels), scale_factor=scale_factor,
                                   device=dev, frames_per_call=frames_per_call)
        eng.load_state_dict(first_stage_model.state_dict())
        object.__setattr__(first_stage_model, "_moca_vae", eng)
    return eng

"""Host-side mirror of the reference first-stage model: decode path, then the encode path (``AutoencoderKLEncoder``).

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


def encoder_param_shapes(ch, ch_mult, num_res_blocks, z_channels, in_channels=3, double_z=True, embed_dim=None):
    """{state_dict key: shape} of Encoder + quant_conv, in the reference's registration order (ae_modules.py:378-428)."""
    embed_dim = z_channels if embed_dim is None else embed_dim
    mom = (2 if double_z else 1) * z_channels
    shapes = {}

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
    conv("encoder.conv_in", ch, in_channels, 3)
    block_in = ch
    for lvl in range(L):
        block_out = ch * ch_mult[lvl]
        for j in range(num_res_blocks):
            res(f"encoder.down.{lvl}.block.{j}", block_in, block_out)
            block_in = block_out
        if lvl != L - 1:
            conv(f"encoder.down.{lvl}.downsample.conv", block_in, block_in, 3)
    res("encoder.mid.block_1", block_in, block_in)
    norm("encoder.mid.attn_1.norm", block_in)
    for n in ("q", "k", "v", "proj_out"):
        conv(f"encoder.mid.attn_1.{n}", block_in, block_in, 1)
    res("encoder.mid.block_2", block_in, block_in)
    norm("encoder.norm_out", block_in)
    conv("encoder.conv_out", mom, block_in, 3)
    shapes["quant_conv.weight"] = (2 * embed_dim, mom, 1, 1)
    shapes["quant_conv.bias"] = (2 * embed_dim,)
    return shapes


class _VaeHandle:
    """What the decoder and the encoder mirrors share: the library handle, the loader and the accounting calls."""
    _h = None
    _loaded = False
    device = None
    shapes = None

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
        """Takes ``AutoencoderKL.state_dict()`` (or ``first_stage_model.*``-stripped checkpoint keys).  Keys outside this object's
        path (the other half of the autoencoder, loss.*) are skipped; a missing key of the path raises."""
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


def _check_ddconfig(who, dd):
    for k, want in _UNSUPPORTED.items():
        if k in dd and dd[k] != want:
            raise _lib.MocaError(f"{who}: ddconfig option {k}={dd[k]!r} is not supported (needs {want!r})")
    if dd.get("attn_resolutions"):
        raise _lib.MocaError(f"{who}: attention inside the resolution levels (attn_resolutions) is not supported")


class AutoencoderKLDecoder(_VaeHandle):
    def __init__(self, ddconfig=None, embed_dim=4, scale_factor=0.18215, device=None, frames_per_call=2, _listing_only=False,
                 **ignored):
        dd = dict(ddconfig or {})
        _check_ddconfig("AutoencoderKLDecoder", dd)
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


class AutoencoderKLEncoder(_VaeHandle):
    """``AutoencoderKL.encode`` (autoencoder.py:97-102): Encoder (ae_modules.py:364-463) + quant_conv -> the posterior's moments."""

    def __init__(self, ddconfig=None, embed_dim=4, scale_factor=0.18215, device=None, frames_per_call=2, _listing_only=False, **ignored):
        dd = dict(ddconfig or {})
        _check_ddconfig("AutoencoderKLEncoder", dd)
        self.ch = int(dd.get("ch", 128))
        self.ch_mult = tuple(int(v) for v in dd.get("ch_mult", (1, 2, 4, 4)))
        self.num_res_blocks = int(dd.get("num_res_blocks", 2))
        self.z_channels = int(dd.get("z_channels", 4))
        self.in_channels = int(dd.get("in_channels", 3))
        self.double_z = bool(dd.get("double_z", True))
        if int(embed_dim) != self.z_channels or not self.double_z:
            raise _lib.MocaError(f"AutoencoderKLEncoder: needs double_z and embed_dim == z_channels (got {self.double_z}, {embed_dim}, "
                                 f"{self.z_channels})")
        self.scale_factor = float(scale_factor)
        self.frames_per_call = int(frames_per_call)
        self.shapes = encoder_param_shapes(self.ch, self.ch_mult, self.num_res_blocks, self.z_channels, self.in_channels, self.double_z)
        self.downscale = 2 ** (len(self.ch_mult) - 1)
        self._h = C.c_void_p()
        self.device = None
        if _listing_only:
            c = None
        else:
            self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
            c = ctx(self.device.index)
        mult = host_i32(self.ch_mult)
        check(lib().moca_vae_encoder_create(c, self.ch, C.cast(mult, C.c_void_p), len(self.ch_mult), self.num_res_blocks,
                                            self.in_channels, self.z_channels, int(self.double_z), C.byref(self._h)))
        self._loaded = False

    @torch.no_grad()
    def encode_moments(self, x):
        """x [n, 3, H, W] in [-1, 1] -> moments [n, 2 zc, H/8, W/8] (mean | logvar), the argument of DiagonalGaussianDistribution."""
        if not self._loaded:
            raise _lib.MocaError("AutoencoderKLEncoder.encode: weights not loaded")
        if x.dim() != 4 or x.shape[1] != self.in_channels:
            raise _lib.MocaError(f"AutoencoderKLEncoder.encode: x must be [n, {self.in_channels}, H, W], got {tuple(x.shape)}")
        n, _, h, w = x.shape
        if h % self.downscale or w % self.downscale:
            raise _lib.MocaError(f"AutoencoderKLEncoder.encode: H, W must be multiples of {self.downscale}, got {h} x {w}")
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        out = torch.empty((n, 2 * self.z_channels, h // self.downscale, w // self.downscale), device=self.device, dtype=torch.float32)
        step = max(1, min(self.frames_per_call, 16))
        with torch.cuda.device(self.device):
            for i in range(0, n, step):
                m = min(step, n - i)
                check(lib().moca_vae_encode(self._h, ptr(x[i:i + m]), m, h, w, ptr(out[i:i + m]), stream_ptr()))
        return out

    @torch.no_grad()
    def encode(self, x, **kwargs):
        """AutoencoderKL.encode: returns a posterior with the reference's ``sample`` / ``mode`` (lvdm/distributions.py:24-43)."""
        return DiagonalGaussian(self.encode_moments(x))

    @torch.no_grad()
    def encode_first_stage_2DAE(self, x, noise=None):
        """LatentDiffusion.encode_first_stage_2DAE (ddpm3d.py:485-490): x [b, 3, t, H, W] -> scale_factor * sample(posterior),
        [b, zc, t, H/8, W/8].  ``noise`` [b, zc, t, h, w] replaces the per-frame host-side randn of the reference's ``sample``."""
        b, c, t, h, w = x.shape
        mom = self.encode_moments(x.permute(0, 2, 1, 3, 4).reshape(b * t, c, h, w))
        mom = mom.reshape(b, t, *mom.shape[1:]).permute(0, 2, 1, 3, 4)
        outs = []
        for i in range(t):
            post = DiagonalGaussian(mom[:, :, i])
            outs.append((self.scale_factor * post.sample(None if noise is None else noise[:, :, i])).unsqueeze(2))
        return torch.cat(outs, dim=2)


class DiagonalGaussian:
    """lvdm/distributions.py:24-43 (the parts inference uses): clamp(logvar, -30, 20), std = exp(logvar / 2); ``sample`` draws its
    noise with the HOST generator, as the reference does (torch.randn(shape) without a device), so seeds reproduce."""

    def __init__(self, parameters):
        self.parameters = parameters
        self.mean, logvar = torch.chunk(parameters, 2, dim=1)
        self.logvar = torch.clamp(logvar, -30.0, 20.0)
        self.std = torch.exp(0.5 * self.logvar)
        self.var = torch.exp(self.logvar)

    def sample(self, noise=None):
        if noise is None:
            noise = torch.randn(self.mean.shape)
        return self.mean + self.std * noise.to(device=self.parameters.device)

    def mode(self):
        return self.mean


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


def attach_vae_encoder(first_stage_model, scale_factor=0.18215, device=None, frames_per_call=2):
    """Builds (once) the library encoder of a reference ``AutoencoderKL`` from its own modules and ``state_dict``."""
    eng = getattr(first_stage_model, "_moca_vae_enc", None)
    if eng is None:
        e = first_stage_model.encoder
        if any(len(d.attn) for d in e.down):
            raise _lib.MocaError("attach_vae_encoder: attention inside the down levels is not supported")
        ch = int(e.ch)
        mult = tuple(int(d.block[0].conv1.out_channels) // ch for d in e.down)
        mom = int(e.conv_out.out_channels)
        zc = int(first_stage_model.quant_conv.out_channels) // 2
        dd = dict(ch=ch, ch_mult=mult, num_res_blocks=int(e.num_res_blocks), z_channels=mom // 2, in_channels=int(e.conv_in.in_channels),
                  double_z=True)
        dev = device if device is not None else next(first_stage_model.parameters()).device
        eng = AutoencoderKLEncoder(dd, embed_dim=zc, scale_factor=scale_factor, device=dev, frames_per_call=frames_per_call)
        eng.load_state_dict(first_stage_model.state_dict())
        object.__setattr__(first_stage_model, "_moca_vae_enc", eng)
    return eng

"""Host-side mirror of the reference ``UNetModel`` (lvdm/modules/networks/openaimodel3d.py:279-578).

``UNet3D`` keeps the reference's constructor vocabulary (the yaml's ``unet_config.params``), its ``state_dict`` keys
(``load_state_dict`` takes the reference checkpoint's UNet tensors unchanged) and its
``forward(x, timesteps, context=None, features_adapter=None, fps=16, **kwargs)`` signature, and runs the whole forward
inside libmoca_b200.so.  It owns no torch parameters: weights live in the library as packed bf16.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import check, ctx, host_i32, lib, ptr, stream_ptr
from .arch import UNetConfig, param_shapes


class UNet3D:
    def __init__(self, in_channels=4, out_channels=4, model_channels=320, attention_resolutions=(4, 2, 1), num_res_blocks=2,
                 channel_mult=(1, 2, 4, 4), num_head_channels=64, context_dim=1024, init_attn_heads=8, device=None,
                 _listing_only=False, **ignored):
        # options the shipped yaml fixes (use_linear, temporal_conv, temporal_attention, temporal_selfatt_only,
        # addition_attention, fps_cond, use_relative_position=False ...) are the only ones implemented; they arrive in
        # **ignored and are checked here so a different config fails loudly instead of silently computing something else
        for k, want in dict(use_linear=True, temporal_conv=True, temporal_attention=True, temporal_selfatt_only=True,
                            use_relative_position=False, use_causal_attention=False, addition_attention=True,
                            fps_cond=True, image_cross_attention=False, use_scale_shift_norm=False, resblock_updown=False,
                            conv_resample=True, dims=2, transformer_depth=1).items():
            if k in ignored and ignored[k] != want:
                raise _lib.MocaError(f"UNet3D: option {k}={ignored[k]!r} is not supported by this build (needs {want!r})")
        self.cfg = UNetConfig(in_channels, out_channels, model_channels, tuple(attention_resolutions), num_res_blocks,
                              tuple(channel_mult), num_head_channels, context_dim, init_attn_heads)
        self.shapes = param_shapes(self.cfg)
        self._h = C.c_void_p()
        self.device = None
        if _listing_only:
            c = None
        else:
            self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
            c = ctx(self.device.index)
        mult = host_i32(self.cfg.channel_mult)
        ar = host_i32(self.cfg.attention_resolutions)
        check(lib().moca_unet_create(c, in_channels, out_channels, model_channels, num_res_blocks, C.cast(mult, C.c_void_p),
                                     len(self.cfg.channel_mult), C.cast(ar, C.c_void_p), len(self.cfg.attention_resolutions),
                                     num_head_channels, context_dim, init_attn_heads, C.byref(self._h)))
        self._loaded = False

    # ---- reference-compatible weight interface ------------------------------------------------------------------
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
        missing = [k for k in self.shapes if k not in sd]
        unexpected = [k for k in sd if k not in self.shapes]
        if strict and (missing or unexpected):
            raise _lib.MocaError(f"load_state_dict: missing {missing[:4]} unexpected {unexpected[:4]}")
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
            # one packing launch for the whole state_dict (the fp32 copies stay alive until the call returns)
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

    # ---- forward ------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, x, timesteps, context=None, features_adapter=None, fps=16, ctx_len=None, **kwargs):
        """UNetModel.forward (openaimodel3d.py:534-578).  x [b,c,t,h,w]; timesteps [b] or, FIFO mode, [b*t] per-frame
        values (the reference infers ``is_fifo`` from x.shape[0] != timesteps.shape[0], :535, which only works for b=1;
        here any [b*t] table is accepted); context [b,L,ctx_dim]; fps int or [b].  ``ctx_len`` [b] marks how many context
        tokens of each sample are real when samples with different prompt lengths are batched.
        ``features_adapter`` and other kwargs (the reference forwards stray sampler kwargs here) are ignored."""
        if features_adapter is not None:
            raise _lib.MocaError("UNet3D.forward: features_adapter is not supported")
        if not self._loaded:
            raise _lib.MocaError("UNet3D.forward: weights not loaded")
        b, c, t, h, w = x.shape
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        ts = [int(v) for v in (timesteps.tolist() if isinstance(timesteps, torch.Tensor) else timesteps)]
        if len(ts) == b and b * t != b:
            ts = [v for v in ts for _ in range(t)]          # repeat_interleave(t) of the per-sample embedding (:545-549)
        elif len(ts) != b * t:
            raise _lib.MocaError(f"UNet3D.forward: {len(ts)} timesteps for b={b}, t={t}")
        if context is None:
            raise _lib.MocaError("UNet3D.forward: context is required")
        context = context.to(device=self.device, dtype=torch.float32).contiguous()
        L = context.shape[1]
        lens = [L] * b if ctx_len is None else [int(v) for v in ctx_len]
        if isinstance(fps, torch.Tensor):
            fl = [int(v) for v in fps.reshape(-1).tolist()]
            if len(fl) == 1:
                fl = fl * b
            elif len(fl) != b:
                raise _lib.MocaError(f"UNet3D.forward: fps has {len(fl)} entries for batch {b} (1 or {b} expected)")
        else:
            fl = [int(fps)] * b
        y = torch.empty((b, self.cfg.out_channels, t, h, w), device=self.device, dtype=torch.float32)
        ts_h, len_h, fps_h = host_i32(ts), host_i32(lens), host_i32(fl)
        with torch.cuda.device(self.device):
            check(lib().moca_unet_forward(self._h, ptr(x), C.cast(ts_h, C.c_void_p), ptr(context), C.cast(len_h, C.c_void_p), L,
                                          C.cast(fps_h, C.c_void_p), b, t, h, w, ptr(y), stream_ptr()))
        return y

    __call__ = forward

    # ---- per-module entry points (moca_unet_layer_forward) ---------------------------------------------------------
    def layer_channels(self, name):
        ci, co = C.c_int32(), C.c_int32()
        check(lib().moca_unet_layer_channels(self._h, name.encode(), C.byref(ci), C.byref(co)))
        return ci.value, co.value

    def layer_forward(self, name, x, emb=None, context=None, ctx_len=None):
        """One ResBlock (+ TemporalConvBlock), SpatialTransformer or TemporalTransformer of the loaded network, by its
        state_dict prefix (openaimodel3d.py:208-234, 269-276; attention.py:262-278, 331-373).  x: [b, c, t, h, w] fp32;
        emb: [b*t, 4*model_channels] (ResBlock); context: [b, L, context_dim] (SpatialTransformer).  Returns [b, c_out, t, h, w]."""
        if not self._loaded:
            raise _lib.MocaError("UNet3D.layer_forward: weights not loaded")
        b, c, t, h, w = x.shape
        ci, co = self.layer_channels(name)
        if c != ci:
            raise _lib.MocaError(f"UNet3D.layer_forward: {name} expects {ci} channels, got {c}")
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        y = torch.empty((b, co, t, h, w), device=self.device, dtype=torch.float32)
        L, len_h = 0, None
        if emb is not None:
            emb = emb.to(device=self.device, dtype=torch.float32).contiguous()
            if tuple(emb.shape) != (b * t, 4 * self.cfg.model_channels):
                raise _lib.MocaError(f"UNet3D.layer_forward: emb must be [{b * t}, {4 * self.cfg.model_channels}], got {tuple(emb.shape)}")
        if context is not None:
            context = context.to(device=self.device, dtype=torch.float32).contiguous()
            if context.shape[0] != b:
                raise _lib.MocaError(f"UNet3D.layer_forward: context batch {context.shape[0]} != {b}")
            L = context.shape[1]
            len_h = host_i32([L] * b if ctx_len is None else [int(v) for v in ctx_len])
        with torch.cuda.device(self.device):
            check(lib().moca_unet_layer_forward(self._h, name.encode(), ptr(x), ptr(emb), ptr(context),
                                                C.cast(len_h, C.c_void_p) if len_h is not None else None, L, b, t, h, w, ptr(y),
                                                stream_ptr()))
        return y

    # ---- parity taps / accounting -------------------------------------------------------------------------------
    def set_taps(self, enable=True):
        check(lib().moca_unet_set_taps(self._h, int(enable)))

    def taps(self):
        out = {}
        n = lib().moca_unet_num_taps(self._h)
        buf = C.create_string_buffer(256)
        rows, ch = C.c_int64(), C.c_int64()
        for i in range(n):
            check(lib().moca_unet_tap_info(self._h, i, buf, 256, C.byref(rows), C.byref(ch)))
            t = torch.empty((rows.value, ch.value), device=self.device, dtype=torch.float32)
            check(lib().moca_unet_tap_read(self._h, i, ptr(t), stream_ptr()))
            out[buf.value.decode()] = t
        return out

    def profile(self, iters=5):
        """Per-op device times of the last planned forward: list of (us, gflop, label)."""
        need = lib().moca_unet_profile(self._h, 1, None, 0, stream_ptr())
        buf = C.create_string_buffer(int(need) + 4096)
        lib().moca_unet_profile(self._h, int(iters), buf, len(buf), stream_ptr())
        rows = []
        for line in buf.value.decode().splitlines():
            i, us, gf, label = line.split(",", 3)
            rows.append((float(us), float(gf), label))
        return rows

    def launches(self):
        return int(lib().moca_unet_launches(self._h))

    def flops(self):
        return float(lib().moca_unet_flops(self._h))

    def workspace_bytes(self):
        return int(lib().moca_unet_workspace_bytes(self._h))

    def weight_bytes(self):
        return int(lib().moca_unet_weight_bytes(self._h))

    def __del__(self):
        try:
            if self._h:
                lib().moca_unet_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

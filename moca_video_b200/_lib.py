"""ctypes binding of libmoca_b200.so (include/moca_b200.h).  There is no fallback: if the library is missing or a
call fails the shim raises; nothing here computes anything on the host."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmoca_b200.so")

_lib = None

i32, i64, f32, f64, vp, cp = C.c_int32, C.c_int64, C.c_float, C.c_double, C.c_void_p, C.c_char_p
pi32, pi64 = C.POINTER(C.c_int32), C.POINTER(C.c_int64)

# name -> (restype, argtypes); mirrors include/moca_b200.h one to one (tests/test_abi.py checks the symbols)
SIGNATURES = {
    "moca_last_error": (cp, []),
    "moca_abi_version": (i32, []),
    "moca_operand_dtype": (i32, []),
    "moca_ctx_create": (i32, [i32, C.POINTER(vp)]),
    "moca_ctx_destroy": (i32, [vp]),
    "moca_cfg_ddim_step": (i32, [vp, vp, vp, vp, vp, f32, vp, vp, vp, i32, vp, vp, f64, f32, f32, i32, f32,
                                 i32, i32, i32, i32, i32, vp, vp, vp]),
    "moca_latent_to_u8": (i32, [vp, vp, i32, i32, i32, i32, i32, vp, vp]),
    "moca_mask_iou": (i32, [vp, vp, i32, vp, i32, i32, i32, C.POINTER(C.c_float), vp]),
    "moca_sam_blend": (i32, [vp, vp, vp, vp, vp, vp, vp, i32, f64, f32, f32, i32, i32, i32, i32, i32, vp, vp]),
    "moca_p_sample_ddim": (i32, [vp, vp, vp, vp, vp, f32, f32, f32, f32, f32, i32, f32, f32, i64, vp, vp, vp]),
    "moca_renoise": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp]),
    "moca_queue_shift": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]),
    "moca_window_gather": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, vp]),
    "moca_window_scatter": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp]),
    "moca_freeinit_filter": (i32, [vp, i32, i32, f64, f64, i32, i32, i32, vp, vp]),
    "moca_freeinit_work_bytes": (i64, [i32, i32, i32, i32]),
    "moca_freeinit_mix": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, vp, vp, vp]),
    "moca_freeinit_mix_static_max_elems": (i64, []),
    "moca_freeinit_mix_static": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, vp, vp]),
    "moca_unet_create": (i32, [vp, i32, i32, i32, i32, vp, i32, vp, i32, i32, i32, i32, C.POINTER(vp)]),
    "moca_unet_destroy": (i32, [vp]),
    "moca_unet_load": (i32, [vp, cp, vp, i64, vp]),
    "moca_unet_load_many": (i32, [vp, i32, C.POINTER(cp), C.POINTER(vp), pi64, vp]),
    "moca_unet_finalize": (i32, [vp]),
    "moca_unet_forward": (i32, [vp, vp, vp, vp, vp, i32, vp, i32, i32, i32, i32, vp, vp]),
    "moca_unet_layer_forward": (i32, [vp, cp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]),
    "moca_unet_layer_channels": (i32, [vp, cp, pi32, pi32]),
    "moca_vae_create": (i32, [vp, i32, vp, i32, i32, i32, i32, C.POINTER(vp)]),
    "moca_vae_decode": (i32, [vp, vp, i32, i32, i32, f32, vp, vp]),
    "moca_vae_encoder_create": (i32, [vp, i32, vp, i32, i32, i32, i32, i32, C.POINTER(vp)]),
    "moca_vae_encode": (i32, [vp, vp, i32, i32, i32, vp, vp]),
    "moca_text_create": (i32, [vp, i32, i32, i32, i32, i32, i32, C.POINTER(vp)]),
    "moca_text_encode": (i32, [vp, vp, i32, vp, vp]),
    "moca_unet_set_taps": (i32, [vp, i32]),
    "moca_unet_num_taps": (i32, [vp]),
    "moca_unet_tap_info": (i32, [vp, i32, cp, i32, pi64, pi64]),
    "moca_unet_tap_read": (i32, [vp, i32, vp, vp]),
    "moca_unet_launches": (i32, [vp]),
    "moca_unet_flops": (f64, [vp]),
    "moca_unet_param_count": (i32, [vp]),
    "moca_unet_param_info": (i32, [vp, i32, cp, i32, pi64]),
    "moca_unet_workspace_bytes": (i64, [vp]),
    "moca_unet_profile": (i64, [vp, i32, cp, i64, vp]),
    "moca_unet_weight_bytes": (i64, [vp]),
    "moca_gemm_op16": (i32, [vp, vp, vp, vp, i32, i32, i32, vp, i32, vp]),
    "moca_geglu_f32": (i32, [vp, vp, vp, vp, i32, i32, i32, vp, vp]),
    "moca_conv3x3_f32": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp]),
    "moca_tconv_f32": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]),
    "moca_groupnorm_op16": (i32, [vp, vp, vp, vp, i64, i32, i32, f32, i32, vp, vp]),
    "moca_layernorm_op16": (i32, [vp, vp, vp, vp, i64, i32, vp, vp]),
    "moca_attention_op16": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, i32, i32, vp]),
    "moca_comm_unique_id": (i32, [vp]),
    "moca_comm_init": (i32, [vp, vp, i32, i32]),
    "moca_queue_handoff": (i32, [vp, vp, vp, i64, vp]),
    "moca_comm_destroy": (i32, [vp]),
    "moca_comm_size": (i32, [vp]),
    "moca_halfwin_pack": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, vp, vp]),
    "moca_halfwin_unpack": (i32, [vp, vp, i32, vp, i32, i32, i32, i32, i32, i32, vp, vp]),
}


class MocaError(RuntimeError):
    pass


def lib():
    """Loads libmoca_b200.so (built in-tree by ``python -m moca_video_b200.build``).  Raises if it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MocaError(f"{LIB_PATH} is missing: build it with `python -m moca_video_b200.build` "
                            "(there is no CPU or PyTorch fallback for this path)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().moca_last_error()
        raise MocaError(f"libmoca_b200 error {rc}: {msg.decode() if msg else '?'}")


_ctx = {}


def ctx(device_index=None):
    """One library context per device of this process."""
    import torch
    if not torch.cuda.is_available():
        raise MocaError("libmoca_b200 needs a CUDA device (sm_100a); no CPU fallback exists")
    if device_index is None:
        device_index = torch.cuda.current_device()
    if device_index not in _ctx:
        h = vp()
        check(lib().moca_ctx_create(int(device_index), C.byref(h)))
        _ctx[device_index] = h
    return _ctx[device_index]


def stream_ptr():
    import torch
    return vp(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return vp(t.data_ptr()) if t is not None else vp(0)


def host_i32(values):
    arr = (C.c_int32 * len(values))(*[int(v) for v in values])
    return arr


def host_f32(values):
    arr = (C.c_float * len(values))(*[float(v) for v in values])
    return arr

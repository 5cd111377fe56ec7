"""Host-side mirror of the reference text conditioner: ``FrozenOpenCLIPEmbedder`` (lvdm/modules/encoders/condition.py:174-235).

``OpenCLIPTextEncoder`` keeps the reference constructor vocabulary (``arch``, ``layer``, ``max_length``), takes
``FrozenOpenCLIPEmbedder.state_dict()`` (keys ``model.*``; the ``cond_stage_model.`` prefix of a full checkpoint is stripped by the
caller) and runs ``encode_with_transformer`` inside libmoca_b200.so.  Tokenisation (open_clip's BPE, host string work) is not part of
the library: ``forward(text)`` uses ``open_clip.tokenize`` when that package imports, or the ``tokenizer`` callable given at
construction; ``encode_with_transformer(tokens)`` takes ids directly.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import check, ctx, lib, ptr, stream_ptr
from .vae import _VaeHandle

# open_clip model configs of the text towers (open_clip/model_configs/*.json: text_cfg): vocab, context, width, heads, layers
ARCHS = {"ViT-H-14": (49408, 77, 1024, 16, 24), "ViT-L-14": (49408, 77, 768, 12, 12), "ViT-B-32": (49408, 77, 512, 8, 12)}
LAYERS = {"last": 0, "penultimate": 1}


def text_param_shapes(vocab, context_length, width, heads, layers_used):
    """{state_dict key: shape} of the part of ``FrozenOpenCLIPEmbedder.state_dict()`` the path reads."""
    s = {"model.token_embedding.weight": (vocab, width), "model.positional_embedding": (context_length, width)}
    for i in range(layers_used):
        p = f"model.transformer.resblocks.{i}"
        s[p + ".ln_1.weight"] = (width,); s[p + ".ln_1.bias"] = (width,)
        s[p + ".attn.in_proj_weight"] = (3 * width, width); s[p + ".attn.in_proj_bias"] = (3 * width,)
        s[p + ".attn.out_proj.weight"] = (width, width); s[p + ".attn.out_proj.bias"] = (width,)
        s[p + ".ln_2.weight"] = (width,); s[p + ".ln_2.bias"] = (width,)
        s[p + ".mlp.c_fc.weight"] = (4 * width, width); s[p + ".mlp.c_fc.bias"] = (4 * width,)
        s[p + ".mlp.c_proj.weight"] = (width, 4 * width); s[p + ".mlp.c_proj.bias"] = (width,)
    s["model.ln_final.weight"] = (width,); s["model.ln_final.bias"] = (width,)
    return s


class OpenCLIPTextEncoder(_VaeHandle):
    def __init__(self, arch="ViT-H-14", version="laion2b_s32b_b79k", device=None, max_length=77, freeze=True, layer="last",
                 tokenizer=None, text_cfg=None, _listing_only=False):
        if layer not in LAYERS:
            raise _lib.MocaError(f"OpenCLIPTextEncoder: layer {layer!r} (one of {sorted(LAYERS)})")
        if text_cfg is None:
            if arch not in ARCHS:
                raise _lib.MocaError(f"OpenCLIPTextEncoder: arch {arch!r} is not known here (known: {sorted(ARCHS)}; or pass text_cfg)")
            text_cfg = ARCHS[arch]
        self.vocab, self.context_length, self.width, self.heads, self.layers = (int(v) for v in text_cfg)
        if int(max_length) != self.context_length:
            raise _lib.MocaError(f"OpenCLIPTextEncoder: max_length {max_length} != context length {self.context_length}")
        self.layer, self.layer_idx = layer, LAYERS[layer]
        self.layers_used = self.layers - self.layer_idx
        self.max_length = int(max_length)
        self.tokenizer = tokenizer
        self.shapes = text_param_shapes(self.vocab, self.context_length, self.width, self.heads, self.layers_used)
        self._h = C.c_void_p()
        self.device = None
        if _listing_only:
            c = None
        else:
            self.device = torch.device(device if device not in (None, "cuda") else f"cuda:{torch.cuda.current_device()}")
            c = ctx(self.device.index)
        check(lib().moca_text_create(c, self.vocab, self.context_length, self.width, self.heads, self.layers, self.layers_used,
                                     C.byref(self._h)))
        self._loaded = False

    def freeze(self):
        return self

    @torch.no_grad()
    def encode_with_transformer(self, text):
        """condition.py:212-219: token ids [b, 77] (any integer dtype) -> z [b, 77, width] fp32."""
        if not self._loaded:
            raise _lib.MocaError("OpenCLIPTextEncoder: weights not loaded")
        if text.dim() != 2 or text.shape[1] != self.context_length:
            raise _lib.MocaError(f"OpenCLIPTextEncoder: tokens must be [b, {self.context_length}], got {tuple(text.shape)}")
        if int(text.min()) < 0 or int(text.max()) >= self.vocab:
            raise _lib.MocaError(f"OpenCLIPTextEncoder: token id outside 0..{self.vocab - 1}")
        tok = text.to(device=self.device, dtype=torch.int32).contiguous()
        b = tok.shape[0]
        z = torch.empty((b, self.context_length, self.width), device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            for i in range(0, b, 64):
                m = min(64, b - i)
                check(lib().moca_text_encode(self._h, ptr(tok[i:i + m]), m, ptr(z[i:i + m]), stream_ptr()))
        return z

    def forward(self, text):
        """condition.py:205-210: list of strings -> z.  Needs a tokenizer (open_clip.tokenize, or the one given at construction)."""
        tok_fn = self.tokenizer
        if tok_fn is None:
            try:
                import open_clip
                tok_fn = open_clip.tokenize
            except ImportError as e:
                raise _lib.MocaError("OpenCLIPTextEncoder.forward: open_clip is not installed and no tokenizer= was given; "
                                     "call encode_with_transformer(token_ids) instead") from e
        return self.encode_with_transformer(tok_fn(text))

    __call__ = forward

    def encode(self, text):
        return self(text)


def attach_text(embedder, device=None):
    """Builds (once) the library text tower of a reference ``FrozenOpenCLIPEmbedder`` from its own ``model`` and ``state_dict``."""
    eng = getattr(embedder, "_moca_text", None)
    if eng is None:
        m = embedder.model
        blocks = m.transformer.resblocks
        cfg = (int(m.token_embedding.weight.shape[0]), int(m.positional_embedding.shape[0]), int(m.token_embedding.weight.shape[1]),
               int(blocks[0].attn.num_heads), len(blocks))
        dev = device if device is not None else m.positional_embedding.device
        layer = {v: k for k, v in LAYERS.items()}[int(embedder.layer_idx)]
        eng = OpenCLIPTextEncoder(device=dev, max_length=cfg[1], layer=layer, text_cfg=cfg)
        eng.load_state_dict(embedder.state_dict())
        object.__setattr__(embedder, "_moca_text", eng)
    return eng

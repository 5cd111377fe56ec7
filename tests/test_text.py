"""Text tower: FrozenOpenCLIPEmbedder.encode_with_transformer behind moca_text_create / moca_text_encode.

The oracle (oracle/text_ref.py) restates open_clip's published text transformer; open_clip is not installed, so the fixture pins it to
transformers.CLIPTextModel with mapped weights instead (see the oracle's header: parity unpinned against the reference's own
dependency).  Bar for the 16-bit-operand CUDA path: rel-L2 <= 2e-2 against fp32.
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

TOL = 2e-2


def rel(a, b):
    a = a.float(); b = b.float()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _small():
    from moca_video_b200.text import text_param_shapes
    from oracle.make_golden import TEXT_SMALL
    from oracle.text_ref import random_text_state_dict
    vocab, L, W, heads, layers = TEXT_SMALL
    return TEXT_SMALL, random_text_state_dict(text_param_shapes(vocab, L, W, heads, layers), seed=8)


@pytest.mark.parametrize("layer,idx", [("last", 0), ("penultimate", 1)])
def test_text_restatement_matches_fixture(layer, idx):
    from oracle.text_ref import text_encode
    (vocab, L, W, heads, layers), sd = _small()
    g = np.load(os.path.join(GOLDEN, "text_small.npz"))
    z = text_encode(sd, torch.from_numpy(g["tokens"]), heads, layers - idx)
    assert rel(z, torch.from_numpy(g["z_" + layer])) < 1e-5


def test_text_loader_parameter_table():
    """moca_text_create(ctx=NULL) lists what the packer consumes: the ViT-H-14 text tower's keys for 23 of 24 blocks."""
    from moca_video_b200 import MocaError
    from moca_video_b200.text import OpenCLIPTextEncoder, text_param_shapes
    e = OpenCLIPTextEncoder(layer="penultimate", _listing_only=True)
    shapes = text_param_shapes(49408, 77, 1024, 16, 23)
    assert e.expected_params() == {k: int(np.prod(v)) for k, v in shapes.items()}
    assert "model.transformer.resblocks.22.mlp.c_proj.weight" in shapes and "model.transformer.resblocks.23.ln_1.weight" not in shapes
    last = OpenCLIPTextEncoder(layer="last", _listing_only=True)
    assert len(last.expected_params()) == len(shapes) + 12
    with pytest.raises(MocaError):
        OpenCLIPTextEncoder(arch="ViT-bigG-14", _listing_only=True)             # width 1280 / 20 heads: not in the table
    with pytest.raises(MocaError):
        OpenCLIPTextEncoder(layer="pooled", _listing_only=True)
    with pytest.raises(MocaError):
        OpenCLIPTextEncoder(text_cfg=(1000, 77, 96, 2, 2), _listing_only=True)   # head dim 48


@pytest.mark.gpu
@pytest.mark.parametrize("layer", ["last", "penultimate"])
def test_text_small_vs_fixture(layer):
    from moca_video_b200.text import OpenCLIPTextEncoder
    cfg, sd = _small()
    g = np.load(os.path.join(GOLDEN, "text_small.npz"))
    enc = OpenCLIPTextEncoder(device="cuda", layer=layer, text_cfg=cfg).load_state_dict(sd)     # "penultimate" ignores block 3's keys
    z = enc.encode_with_transformer(torch.from_numpy(g["tokens"]))
    assert z.shape == g["z_" + layer].shape
    assert rel(z.cpu(), torch.from_numpy(g["z_" + layer])) < TOL
    # every sample alone gives the same rows (samples are independent; causal mask inside a sample)
    z1 = enc.encode_with_transformer(torch.from_numpy(g["tokens"][1:2]))
    assert rel(z1, z[1:2]) < 4e-3


@pytest.mark.gpu
def test_text_vit_h_vs_oracle():
    """The shipped conditioner: ViT-H-14 text tower, layer = penultimate, two prompts (the CFG pair)."""
    from moca_video_b200.text import OpenCLIPTextEncoder, text_param_shapes
    from oracle.text_ref import random_text_state_dict, text_encode
    torch.backends.cuda.matmul.allow_tf32 = False
    sd = {k: v.cuda() for k, v in random_text_state_dict(text_param_shapes(49408, 77, 1024, 16, 23), seed=9).items()}
    gen = torch.Generator().manual_seed(3)
    tokens = torch.randint(1, 49405, (2, 77), generator=gen)
    tokens[:, 0] = 49406
    tokens[0, 9] = 49407; tokens[0, 10:] = 0
    tokens[1, 1] = 49407; tokens[1, 2:] = 0                                      # the empty prompt of classifier-free guidance
    enc = OpenCLIPTextEncoder(device="cuda", layer="penultimate").load_state_dict(sd)
    z = enc.encode_with_transformer(tokens)
    ref = text_encode(sd, tokens.cuda(), 16, 23)
    assert z.shape == (2, 77, 1024) and torch.isfinite(z).all()
    assert rel(z, ref) < TOL
    z2 = enc.encode_with_transformer(tokens)                                     # graph replay
    assert torch.equal(z, z2)


@pytest.mark.gpu
def test_text_rejects_bad_input():
    from moca_video_b200 import MocaError
    from moca_video_b200.text import OpenCLIPTextEncoder
    cfg, sd = _small()
    enc = OpenCLIPTextEncoder(device="cuda", text_cfg=cfg)
    with pytest.raises(MocaError):
        enc.encode_with_transformer(torch.zeros(1, 77, dtype=torch.long))         # weights not loaded
    enc.load_state_dict(sd)
    with pytest.raises(MocaError):
        enc.encode_with_transformer(torch.zeros(1, 76, dtype=torch.long))
    with pytest.raises(MocaError):
        enc.encode_with_transformer(torch.full((1, 77), 1000, dtype=torch.long))  # id == vocab
    with pytest.raises(MocaError):
        enc(["a cat"])                                                           # no tokenizer here (open_clip absent)
    enc.tokenizer = lambda texts: torch.ones(len(texts), 77, dtype=torch.long)
    assert enc(["a cat", ""]).shape == (2, 77, 128)


class _Block(torch.nn.Module):
    """open_clip's ResidualAttentionBlock as published: the same sub-module names and the same torch ops (nn.MultiheadAttention with
    the additive causal mask, Sequential(c_fc, GELU, c_proj)); open_clip itself is not installed."""

    def __init__(self, W, heads):
        super().__init__()
        from collections import OrderedDict
        nn = torch.nn
        self.ln_1 = nn.LayerNorm(W)
        self.attn = nn.MultiheadAttention(W, heads)
        self.ln_2 = nn.LayerNorm(W)
        self.mlp = nn.Sequential(OrderedDict([("c_fc", nn.Linear(W, 4 * W)), ("gelu", nn.GELU()), ("c_proj", nn.Linear(4 * W, W))]))

    def forward(self, x, attn_mask=None):                     # x: [L, N, D] like open_clip's text path at 2.30 (condition.py:215)
        h = self.ln_1(x)
        x = x + self.attn(h, h, h, need_weights=False, attn_mask=attn_mask)[0]
        return x + self.mlp(self.ln_2(x))


class _Embedder(torch.nn.Module):
    """Stand-in with FrozenOpenCLIPEmbedder's attribute layout (condition.py:188-203) and its own torch forward (:212-230)."""

    def __init__(self, cfg, layer_idx, sd):
        super().__init__()
        nn = torch.nn
        vocab, L, W, heads, layers = cfg
        m = nn.Module()
        m.token_embedding = nn.Embedding(vocab, W)
        m.positional_embedding = nn.Parameter(torch.zeros(L, W))
        m.transformer = nn.Module()
        m.transformer.resblocks = nn.ModuleList([_Block(W, heads) for _ in range(layers)])
        m.ln_final = nn.LayerNorm(W)
        m.text_projection = nn.Parameter(torch.zeros(W, W))   # present in the state_dict, not on the path
        m.logit_scale = nn.Parameter(torch.ones([]))
        m.register_buffer("attn_mask", torch.full((L, L), float("-inf")).triu_(1), persistent=False)
        self.model, self.layer_idx = m, layer_idx
        missing, unexpected = self.load_state_dict(sd, strict=False)
        assert not unexpected and set(missing) <= {"model.text_projection", "model.logit_scale"}, (missing, unexpected)

    def encode_with_transformer(self, text):
        m = self.model
        x = (m.token_embedding(text) + m.positional_embedding).permute(1, 0, 2)
        for i, r in enumerate(m.transformer.resblocks):
            if i == len(m.transformer.resblocks) - self.layer_idx:
                break
            x = r(x, attn_mask=m.attn_mask)
        return m.ln_final(x.permute(1, 0, 2))


@pytest.mark.parametrize("layer_idx", [0, 1])
def test_text_restatement_matches_torch_module_assembly(layer_idx):
    """A second, independent pin: the modules open_clip assembles (nn.MultiheadAttention + causal mask) run by torch itself."""
    from oracle.text_ref import text_encode
    cfg, sd = _small()
    g = np.load(os.path.join(GOLDEN, "text_small.npz"))
    tokens = torch.from_numpy(g["tokens"])
    with torch.no_grad():
        ref = _Embedder(cfg, layer_idx, sd).eval().encode_with_transformer(tokens)
    assert rel(text_encode(sd, tokens, cfg[3], cfg[4] - layer_idx), ref) < 1e-5
    assert rel(ref, torch.from_numpy(g["z_last" if layer_idx == 0 else "z_penultimate"])) < 1e-5


@pytest.mark.gpu
def test_attach_text_reads_the_reference_module_tree():
    from moca_video_b200.patch import _clip_encode_with_transformer
    from moca_video_b200.text import attach_text
    cfg, sd = _small()
    g = np.load(os.path.join(GOLDEN, "text_small.npz"))
    emb = _Embedder(cfg, 1, sd).cuda().eval()
    eng = attach_text(emb)
    assert (eng.vocab, eng.context_length, eng.width, eng.heads, eng.layers, eng.layers_used) == (1000, 77, 128, 2, 4, 3)
    assert attach_text(emb) is eng
    tokens = torch.from_numpy(g["tokens"]).cuda()
    z = _clip_encode_with_transformer(emb, tokens)
    with torch.no_grad():
        ref = emb.encode_with_transformer(tokens)
    assert rel(z, ref) < TOL and rel(z.cpu(), torch.from_numpy(g["z_penultimate"])) < TOL

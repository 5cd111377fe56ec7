"""Plain-PyTorch fp32 restatement of the reference text conditioner, driven by a ``state_dict``.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

PARITY UNPINNED against the reference's own dependency: ``FrozenOpenCLIPEmbedder`` (lvdm/modules/encoders/condition.py:174-235) runs
the text transformer of open_clip_torch==2.30.0 (reference requirements.txt:196), which is neither vendored in /root/reference nor
installed in this image.  This file restates that package's published algorithm (open_clip/transformer.py: ``ResidualAttentionBlock``,
``Transformer``; open_clip/model.py: ``CLIP.__init__`` text part, ``build_attention_mask``) behind the reference's own call site
(condition.py:212-230), and is cross-checked — ``oracle/make_golden.py text_small`` -> ``tests/golden/text_small.npz`` — against
``transformers.CLIPTextModel`` (installed here), an independent implementation of the same published model, with the weights mapped
key by key (``to_hf_state_dict``).  The state_dict keys are open_clip's (as they appear in the VideoCrafter2 checkpoint under
``cond_stage_model.model.*``).

  FrozenOpenCLIPEmbedder.encode_with_transformer   condition.py:212-219
  FrozenOpenCLIPEmbedder.text_transformer_forward  condition.py:221-230   (stops ``layer_idx`` blocks before the end)
  ResidualAttentionBlock.forward                   open_clip/transformer.py: x + attn(ln_1(x)); x + mlp(ln_2(x)), mlp = c_fc, GELU, c_proj
  build_attention_mask                             open_clip/model.py: -inf above the diagonal (causal)
"""
import torch
import torch.nn.functional as F


def text_encode(sd, tokens, heads, layers_used):
    """tokens [b, L] int -> z [b, L, width]: encode_with_transformer with ``layers_used`` residual blocks, then ln_final."""
    sd = {k: v.float() for k, v in sd.items()}
    x = sd["model.token_embedding.weight"][tokens.long()] + sd["model.positional_embedding"]
    b, L, W = x.shape
    d = W // heads
    mask = torch.full((L, L), float("-inf"), device=x.device).triu_(1)
    for i in range(layers_used):
        p = f"model.transformer.resblocks.{i}"
        h = F.layer_norm(x, (W,), sd[p + ".ln_1.weight"], sd[p + ".ln_1.bias"], 1e-5)
        qkv = F.linear(h, sd[p + ".attn.in_proj_weight"], sd[p + ".attn.in_proj_bias"])
        q, k, v = (t.reshape(b, L, heads, d).permute(0, 2, 1, 3) for t in qkv.chunk(3, dim=-1))
        a = torch.softmax(q @ k.transpose(-1, -2) * (d ** -0.5) + mask, dim=-1) @ v
        a = a.permute(0, 2, 1, 3).reshape(b, L, W)
        x = x + F.linear(a, sd[p + ".attn.out_proj.weight"], sd[p + ".attn.out_proj.bias"])
        h = F.layer_norm(x, (W,), sd[p + ".ln_2.weight"], sd[p + ".ln_2.bias"], 1e-5)
        h = F.gelu(F.linear(h, sd[p + ".mlp.c_fc.weight"], sd[p + ".mlp.c_fc.bias"]))
        x = x + F.linear(h, sd[p + ".mlp.c_proj.weight"], sd[p + ".mlp.c_proj.bias"])
    return F.layer_norm(x, (W,), sd["model.ln_final.weight"], sd["model.ln_final.bias"], 1e-5)


def random_text_state_dict(shapes, seed=0):
    """Seeded weights with open_clip's initialisation scales (CLIP.init_parameters): same on every machine."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shp in shapes.items():
        if k.endswith(("ln_1.weight", "ln_2.weight", "ln_final.weight")):
            sd[k] = 1.0 + 0.1 * torch.randn(shp, generator=g)
        elif len(shp) == 1:
            sd[k] = 0.02 * torch.randn(shp, generator=g)
        elif "embedding" in k:
            sd[k] = 0.02 * torch.randn(shp, generator=g) if "token" in k else 0.01 * torch.randn(shp, generator=g)
        else:
            sd[k] = torch.randn(shp, generator=g) * (shp[1] ** -0.5)
    return sd


def to_hf_state_dict(sd, layers):
    """open_clip text-tower keys -> transformers.CLIPTextModel keys (every block present in ``sd``)."""
    out = {"text_model.embeddings.token_embedding.weight": sd["model.token_embedding.weight"],
           "text_model.embeddings.position_embedding.weight": sd["model.positional_embedding"],
           "text_model.final_layer_norm.weight": sd["model.ln_final.weight"],
           "text_model.final_layer_norm.bias": sd["model.ln_final.bias"]}
    for i in range(layers):
        p, q = f"model.transformer.resblocks.{i}", f"text_model.encoder.layers.{i}"
        W = sd[p + ".ln_1.weight"].shape[0]
        for j, n in enumerate(("q_proj", "k_proj", "v_proj")):
            out[f"{q}.self_attn.{n}.weight"] = sd[p + ".attn.in_proj_weight"][j * W:(j + 1) * W]
            out[f"{q}.self_attn.{n}.bias"] = sd[p + ".attn.in_proj_bias"][j * W:(j + 1) * W]
        for a, b in ((".attn.out_proj", ".self_attn.out_proj"), (".ln_1", ".layer_norm1"), (".ln_2", ".layer_norm2"),
                     (".mlp.c_fc", ".mlp.fc1"), (".mlp.c_proj", ".mlp.fc2")):
            out[q + b + ".weight"] = sd[p + a + ".weight"]
            out[q + b + ".bias"] = sd[p + a + ".bias"]
    return out

"""Architecture description of the VideoCrafter2 3D-UNet (host-side logic, no compute).

Mirrors the constructor of the reference ``UNetModel`` (lvdm/modules/networks/openaimodel3d.py:306-532)
for the option set the shipped yaml uses (configs/inference_t2v_512_v2.0.yaml:22-50): it enumerates the
blocks, and from them every ``state_dict`` key and shape, without instantiating any torch module.
The same block list drives the CUDA runner (csrc/unet.cu builds the identical list from the same
integers) and the oracle restatement (oracle/unet_ref.py).
"""
import zlib
from dataclasses import dataclass, field
from typing import List, Tuple

import torch


@dataclass
class UNetConfig:
    in_channels: int = 4
    out_channels: int = 4
    model_channels: int = 320
    attention_resolutions: Tuple[int, ...] = (4, 2, 1)
    num_res_blocks: int = 2
    channel_mult: Tuple[int, ...] = (1, 2, 4, 4)
    num_head_channels: int = 64
    context_dim: int = 1024
    init_attn_heads: int = 8          # openaimodel3d.py:382 (hard-coded n_heads=8)

    @staticmethod
    def videocrafter2():
        return UNetConfig()


@dataclass
class Layer:
    kind: str                 # 'res' | 'st' | 'tt' | 'down' | 'up' | 'conv_in'
    name: str                 # state_dict prefix
    cin: int = 0
    cout: int = 0
    heads: int = 0
    inner: int = 0            # transformer inner dim
    linear: bool = True       # TemporalTransformer proj as nn.Linear (True) or Conv1d k=1 (init_attn)


@dataclass
class Block:
    name: str
    layers: List[Layer] = field(default_factory=list)
    skip_cin: int = 0         # channels popped from the skip stack and concatenated (output blocks)


def build_blocks(cfg: UNetConfig):
    """Returns (input_blocks, init_attn, middle, output_blocks) following openaimodel3d.py:373-526."""
    mc, dh = cfg.model_channels, cfg.num_head_channels
    inp = [Block("input_blocks.0", [Layer("conv_in", "input_blocks.0.0", cfg.in_channels, mc)])]
    init_attn = Block("init_attn", [Layer("tt", "init_attn.0", mc, mc, cfg.init_attn_heads,
                                          cfg.init_attn_heads * dh, linear=False)])
    chans = [mc]
    ch, ds = mc, 1
    bi = 1
    for level, mult in enumerate(cfg.channel_mult):
        for _ in range(cfg.num_res_blocks):
            b = Block(f"input_blocks.{bi}")
            b.layers.append(Layer("res", f"input_blocks.{bi}.0", ch, mult * mc))
            ch = mult * mc
            if ds in cfg.attention_resolutions:
                b.layers.append(Layer("st", f"input_blocks.{bi}.1", ch, ch, ch // dh, ch))
                b.layers.append(Layer("tt", f"input_blocks.{bi}.2", ch, ch, ch // dh, ch))
            inp.append(b); chans.append(ch); bi += 1
        if level != len(cfg.channel_mult) - 1:
            inp.append(Block(f"input_blocks.{bi}", [Layer("down", f"input_blocks.{bi}.0", ch, ch)]))
            chans.append(ch); bi += 1; ds *= 2
    mid = Block("middle_block", [
        Layer("res", "middle_block.0", ch, ch),
        Layer("st", "middle_block.1", ch, ch, ch // dh, ch),
        Layer("tt", "middle_block.2", ch, ch, ch // dh, ch),
        Layer("res", "middle_block.3", ch, ch)])
    out = []
    oi = 0
    for level, mult in list(enumerate(cfg.channel_mult))[::-1]:
        for i in range(cfg.num_res_blocks + 1):
            ich = chans.pop()
            b = Block(f"output_blocks.{oi}", skip_cin=ich)
            b.layers.append(Layer("res", f"output_blocks.{oi}.0", ch + ich, mc * mult))
            ch = mc * mult
            li = 1
            if ds in cfg.attention_resolutions:
                b.layers.append(Layer("st", f"output_blocks.{oi}.1", ch, ch, ch // dh, ch))
                b.layers.append(Layer("tt", f"output_blocks.{oi}.2", ch, ch, ch // dh, ch))
                li = 3
            if level and i == cfg.num_res_blocks:
                b.layers.append(Layer("up", f"output_blocks.{oi}.{li}", ch, ch))
                ds //= 2
            out.append(b); oi += 1
    return inp, init_attn, mid, out


def _tf_shapes(p, C, inner, ctx_dim, linear, shapes):
    """SpatialTransformer / TemporalTransformer parameters (lvdm/modules/attention.py:223-329, 187-201, 43-57, 376-398)."""
    shapes[p + ".norm.weight"] = (C,); shapes[p + ".norm.bias"] = (C,)
    if linear:
        shapes[p + ".proj_in.weight"] = (inner, C)
        shapes[p + ".proj_out.weight"] = (C, inner)
    else:
        shapes[p + ".proj_in.weight"] = (inner, C, 1)
        shapes[p + ".proj_out.weight"] = (C, inner, 1)
    shapes[p + ".proj_in.bias"] = (inner,); shapes[p + ".proj_out.bias"] = (C,)
    t = p + ".transformer_blocks.0"
    for a, cd in (("attn1", inner), ("attn2", ctx_dim)):
        shapes[f"{t}.{a}.to_q.weight"] = (inner, inner)
        shapes[f"{t}.{a}.to_k.weight"] = (inner, cd)
        shapes[f"{t}.{a}.to_v.weight"] = (inner, cd)
        shapes[f"{t}.{a}.to_out.0.weight"] = (inner, inner)
        shapes[f"{t}.{a}.to_out.0.bias"] = (inner,)
    shapes[f"{t}.ff.net.0.proj.weight"] = (8 * inner, inner); shapes[f"{t}.ff.net.0.proj.bias"] = (8 * inner,)
    shapes[f"{t}.ff.net.2.weight"] = (inner, 4 * inner); shapes[f"{t}.ff.net.2.bias"] = (inner,)
    for n in ("norm1", "norm2", "norm3"):
        shapes[f"{t}.{n}.weight"] = (inner,); shapes[f"{t}.{n}.bias"] = (inner,)


def param_shapes(cfg: UNetConfig):
    """Ordered dict name -> shape, equal to ``UNetModel(**yaml).state_dict()`` keys/shapes."""
    mc = cfg.model_channels
    te = 4 * mc
    shapes = {}
    for pre in ("time_embed", "fps_embedding"):
        shapes[f"{pre}.0.weight"] = (te, mc); shapes[f"{pre}.0.bias"] = (te,)
        shapes[f"{pre}.2.weight"] = (te, te); shapes[f"{pre}.2.bias"] = (te,)
    inp, init_attn, mid, out = build_blocks(cfg)
    for blk in inp[:1] + [init_attn] + inp[1:] + [mid] + out:
        for L in blk.layers:
            p = L.name
            if L.kind == "conv_in":
                shapes[p + ".weight"] = (L.cout, L.cin, 3, 3); shapes[p + ".bias"] = (L.cout,)
            elif L.kind == "res":
                shapes[p + ".in_layers.0.weight"] = (L.cin,); shapes[p + ".in_layers.0.bias"] = (L.cin,)
                shapes[p + ".in_layers.2.weight"] = (L.cout, L.cin, 3, 3); shapes[p + ".in_layers.2.bias"] = (L.cout,)
                shapes[p + ".emb_layers.1.weight"] = (L.cout, te); shapes[p + ".emb_layers.1.bias"] = (L.cout,)
                shapes[p + ".out_layers.0.weight"] = (L.cout,); shapes[p + ".out_layers.0.bias"] = (L.cout,)
                shapes[p + ".out_layers.3.weight"] = (L.cout, L.cout, 3, 3); shapes[p + ".out_layers.3.bias"] = (L.cout,)
                if L.cin != L.cout:
                    shapes[p + ".skip_connection.weight"] = (L.cout, L.cin, 1, 1)
                    shapes[p + ".skip_connection.bias"] = (L.cout,)
                for k, gi in (("conv1", 0), ("conv2", 0), ("conv3", 0), ("conv4", 0)):
                    ci = 2 if k == "conv1" else 3          # nn.Sequential index of the Conv3d (openaimodel3d.py:252-263)
                    q = f"{p}.temopral_conv.{k}"
                    shapes[f"{q}.0.weight"] = (L.cout,); shapes[f"{q}.0.bias"] = (L.cout,)
                    shapes[f"{q}.{ci}.weight"] = (L.cout, L.cout, 3, 1, 1); shapes[f"{q}.{ci}.bias"] = (L.cout,)
            elif L.kind == "st":
                _tf_shapes(p, L.cin, L.inner, cfg.context_dim, True, shapes)
            elif L.kind == "tt":
                # temporal_selfatt_only => context_dim None => attn2 is self-attention (attention.py:313-314)
                _tf_shapes(p, L.cin, L.inner, L.inner, L.linear, shapes)
            elif L.kind == "down":
                shapes[p + ".op.weight"] = (L.cout, L.cin, 3, 3); shapes[p + ".op.bias"] = (L.cout,)
            elif L.kind == "up":
                shapes[p + ".conv.weight"] = (L.cout, L.cin, 3, 3); shapes[p + ".conv.bias"] = (L.cout,)
    shapes["out.0.weight"] = (mc,); shapes["out.0.bias"] = (mc,)
    shapes["out.2.weight"] = (cfg.out_channels, mc, 3, 3); shapes["out.2.bias"] = (cfg.out_channels,)
    return shapes


def random_state_dict(cfg: UNetConfig, seed=0, device="cpu", dtype=torch.float32):
    """Seeded synthetic weights ("random-init weights of that architecture", BASELINE.json).

    Every tensor is drawn from its own generator (seed ^ crc32(name)), so the result does not depend on
    module construction order.  Nothing is left at zero (the reference's zero-initialised output
    convs make a default-init net output exactly 0).  Norm scales ~ 1 + 0.1 N, biases 0.02 N,
    matrices N(0, 1/fan_in).
    """
    sd = {}
    dev = torch.device(device)
    for name, shp in param_shapes(cfg).items():
        g = torch.Generator(device=dev).manual_seed((seed * 1000003) ^ zlib.crc32(name.encode()))
        r = torch.randn(shp, generator=g, device=dev, dtype=torch.float32)
        if len(shp) == 1:
            is_norm = name.endswith("weight")
            t = 1.0 + 0.1 * r if is_norm else 0.02 * r
        else:
            fan_in = 1
            for d in shp[1:]:
                fan_in *= d
            t = r * (1.0 / fan_in) ** 0.5
        sd[name] = t.to(dtype)
    return sd

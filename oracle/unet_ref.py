"""Plain-PyTorch fp32 restatement of the reference UNet forward, driven by a ``state_dict``.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  It is the floating-point reference the CUDA
path is compared against on the GPU box (where ``/root/reference`` does not exist) and the CPU
baseline that ``bench.py`` times.  It is pinned to the unmodified reference by
``tests/golden/unet_small.npz`` (reference ``UNetModel`` output on seeded weights, produced by
``oracle/make_golden.py``).

Follows, function by function:
  UNetModel.forward            lvdm/modules/networks/openaimodel3d.py:534-578
  TimestepEmbedSequential      openaimodel3d.py:36-47
  ResBlock._forward            openaimodel3d.py:208-234
  TemporalConvBlock.forward    openaimodel3d.py:269-276
  Downsample / Upsample        openaimodel3d.py:75-77, 98-106
  SpatialTransformer.forward   lvdm/modules/attention.py:262-278
  TemporalTransformer.forward  attention.py:331-373
  BasicTransformerBlock        attention.py:216-220
  CrossAttention.forward       attention.py:76-127 (einsum path; xformers path is mathematically equal)
  GEGLU / FeedForward          attention.py:376-403
  timestep_embedding           lvdm/models/utils_diffusion.py:8-28
  GroupNormSpecific            lvdm/basics.py:76-87 (fp32 GroupNorm, eps 1e-5)
"""
import math

import torch
import torch.nn.functional as F

from moca_video_b200.arch import UNetConfig, build_blocks


def timestep_embedding(t, dim, max_period=10000):
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half).to(t.device)
    args = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


def _gn(x, sd, p, eps):
    return F.group_norm(x.float(), 32, sd[p + ".weight"], sd[p + ".bias"], eps)


def _attn(x, ctx, sd, p, heads):
    """CrossAttention.forward (attention.py:76-127), no mask / relative position / image tokens."""
    q = F.linear(x, sd[p + ".to_q.weight"])
    c = x if ctx is None else ctx
    k = F.linear(c, sd[p + ".to_k.weight"])
    v = F.linear(c, sd[p + ".to_v.weight"])
    B, N, C = q.shape
    d = C // heads
    sp = lambda t: t.reshape(B, -1, heads, d).permute(0, 2, 1, 3).reshape(B * heads, -1, d)
    q, k, v = sp(q), sp(k), sp(v)
    out = torch.empty_like(q)
    step = max(1, (1 << 28) // max(1, N * k.shape[1]))      # chunk to bound the score matrix
    for s in range(0, q.shape[0], step):
        sim = torch.einsum("bid,bjd->bij", q[s:s + step], k[s:s + step]) * (d ** -0.5)
        out[s:s + step] = torch.einsum("bij,bjd->bid", sim.softmax(dim=-1), v[s:s + step])
    out = out.reshape(B, heads, N, d).permute(0, 2, 1, 3).reshape(B, N, C)
    return F.linear(out, sd[p + ".to_out.0.weight"], sd[p + ".to_out.0.bias"])


def _tblock(x, ctx, sd, p, heads):
    """BasicTransformerBlock._forward (attention.py:216-220)."""
    ln = lambda t, n: F.layer_norm(t, (t.shape[-1],), sd[f"{p}.{n}.weight"], sd[f"{p}.{n}.bias"], 1e-5)
    x = _attn(ln(x, "norm1"), None, sd, p + ".attn1", heads) + x
    x = _attn(ln(x, "norm2"), ctx, sd, p + ".attn2", heads) + x
    h = F.linear(ln(x, "norm3"), sd[p + ".ff.net.0.proj.weight"], sd[p + ".ff.net.0.proj.bias"])
    a, gate = h.chunk(2, dim=-1)
    h = a * F.gelu(gate)
    return F.linear(h, sd[p + ".ff.net.2.weight"], sd[p + ".ff.net.2.bias"]) + x


def spatial_transformer(x, ctx, sd, p, heads):
    n, c, h, w = x.shape
    y = _gn(x, sd, p + ".norm", 1e-6)
    y = y.permute(0, 2, 3, 1).reshape(n, h * w, c)
    y = F.linear(y, sd[p + ".proj_in.weight"], sd[p + ".proj_in.bias"])
    y = _tblock(y, ctx, sd, p + ".transformer_blocks.0", heads)
    y = F.linear(y, sd[p + ".proj_out.weight"], sd[p + ".proj_out.bias"])
    return y.reshape(n, h, w, c).permute(0, 3, 1, 2) + x


def temporal_transformer(x, sd, p, heads):
    """x: [b,c,t,h,w]; both attentions are self-attention over t (temporal_selfatt_only)."""
    b, c, t, h, w = x.shape
    y = _gn(x, sd, p + ".norm", 1e-6)                        # 5-D input: statistics span all frames
    y = y.permute(0, 3, 4, 2, 1).reshape(b * h * w, t, c)    # (b h w) t c
    wi, wo = sd[p + ".proj_in.weight"], sd[p + ".proj_out.weight"]
    y = F.linear(y, wi.reshape(wi.shape[0], wi.shape[1]), sd[p + ".proj_in.bias"])   # Linear or Conv1d(k=1)
    y = _tblock(y, None, sd, p + ".transformer_blocks.0", heads)
    y = F.linear(y, wo.reshape(wo.shape[0], wo.shape[1]), sd[p + ".proj_out.bias"])
    return y.reshape(b, h, w, t, c).permute(0, 4, 3, 1, 2) + x


def temporal_conv_block(x, sd, p):
    """x: [b,c,t,h,w] (openaimodel3d.py:269-276)."""
    y = x
    for k in ("conv1", "conv2", "conv3", "conv4"):
        ci = 2 if k == "conv1" else 3
        y = F.silu(F.group_norm(y, 32, sd[f"{p}.{k}.0.weight"], sd[f"{p}.{k}.0.bias"], 1e-5))
        y = F.conv3d(y, sd[f"{p}.{k}.{ci}.weight"], sd[f"{p}.{k}.{ci}.bias"], padding=(1, 0, 0))
    return y + x


def resblock(x, emb, sd, p, b):
    h = F.conv2d(F.silu(_gn(x, sd, p + ".in_layers.0", 1e-5)), sd[p + ".in_layers.2.weight"],
                 sd[p + ".in_layers.2.bias"], padding=1)
    e = F.linear(F.silu(emb), sd[p + ".emb_layers.1.weight"], sd[p + ".emb_layers.1.bias"])
    h = h + e[:, :, None, None]
    h = F.conv2d(F.silu(_gn(h, sd, p + ".out_layers.0", 1e-5)), sd[p + ".out_layers.3.weight"],
                 sd[p + ".out_layers.3.bias"], padding=1)
    if (p + ".skip_connection.weight") in sd:
        x = F.conv2d(x, sd[p + ".skip_connection.weight"], sd[p + ".skip_connection.bias"])
    h = x + h
    n, c, hh, ww = h.shape
    h5 = h.reshape(b, n // b, c, hh, ww).permute(0, 2, 1, 3, 4)
    h5 = temporal_conv_block(h5, sd, p + ".temopral_conv")
    return h5.permute(0, 2, 1, 3, 4).reshape(n, c, hh, ww)


def _run_block(blk, h, emb, ctx, sd, b, taps=None):
    for L in blk.layers:
        if L.kind == "conv_in":
            h = F.conv2d(h, sd[L.name + ".weight"], sd[L.name + ".bias"], padding=1)
        elif L.kind == "res":
            h = resblock(h, emb, sd, L.name, b)
        elif L.kind == "st":
            h = spatial_transformer(h, ctx, sd, L.name, L.heads)
        elif L.kind == "tt":
            n, c, hh, ww = h.shape
            h5 = h.reshape(b, n // b, c, hh, ww).permute(0, 2, 1, 3, 4)
            h5 = temporal_transformer(h5, sd, L.name, L.heads)
            h = h5.permute(0, 2, 1, 3, 4).reshape(n, c, hh, ww)
        elif L.kind == "down":
            h = F.conv2d(h, sd[L.name + ".op.weight"], sd[L.name + ".op.bias"], stride=2, padding=1)
        elif L.kind == "up":
            h = F.interpolate(h, scale_factor=2, mode="nearest")
            h = F.conv2d(h, sd[L.name + ".conv.weight"], sd[L.name + ".conv.bias"], padding=1)
        if taps is not None:
            taps[L.name] = h.permute(0, 2, 3, 1).reshape(-1, h.shape[1]).clone()     # channels-last tokens [rows, C]
    return h


@torch.no_grad()
def unet_forward(sd, cfg: UNetConfig, x, timesteps, context, fps=16, taps=None):
    """x [b,c,t,h,w]; timesteps [b] or [b*t] (``is_fifo``, openaimodel3d.py:535); context [b,L,ctx_dim]."""
    mc = cfg.model_channels
    b, _, t, _, _ = x.shape
    is_fifo = x.shape[0] != timesteps.shape[0]
    lin = lambda v, p: F.linear(v, sd[p + ".weight"], sd[p + ".bias"])
    emb = lin(F.silu(lin(timestep_embedding(timesteps, mc), "time_embed.0")), "time_embed.2")
    if isinstance(fps, int):
        fps = torch.full_like(timesteps, fps)
    emb = emb + lin(F.silu(lin(timestep_embedding(fps, mc), "fps_embedding.0")), "fps_embedding.2")
    ctx = context.repeat_interleave(repeats=t, dim=0)
    if not is_fifo:
        emb = emb.repeat_interleave(repeats=t, dim=0)
    h = x.permute(0, 2, 1, 3, 4).reshape(b * t, x.shape[1], x.shape[3], x.shape[4]).float()
    inp, init_attn, mid, out = build_blocks(cfg)
    hs = []
    for i, blk in enumerate(inp):
        h = _run_block(blk, h, emb, ctx, sd, b, taps)
        if i == 0:
            h = _run_block(init_attn, h, emb, ctx, sd, b, taps)
        hs.append(h)
    h = _run_block(mid, h, emb, ctx, sd, b, taps)
    for blk in out:
        h = torch.cat([h, hs.pop()], dim=1)
        h = _run_block(blk, h, emb, ctx, sd, b, taps)
    y = F.conv2d(F.silu(_gn(h, sd, "out.0", 1e-5)), sd["out.2.weight"], sd["out.2.bias"], padding=1)
    n, c, hh, ww = y.shape
    return y.reshape(b, t, c, hh, ww).permute(0, 2, 1, 3, 4).contiguous()

"""Plain-PyTorch fp32 restatement of the reference first-stage decode path, driven by a ``state_dict``.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Pinned to the unmodified reference by ``tests/golden/vae_small.npz``
(reference ``Decoder`` + ``post_quant_conv`` output on seeded weights, produced by ``oracle/make_golden.py vae_small``).

Follows, function by function:
  LatentDiffusion.decode_first_stage_2DAE   lvdm/models/ddpm3d.py:555-562   (1 / scale_factor, per-frame loop)
  AutoencoderKL.decode                      lvdm/models/autoencoder.py:104-107
  Decoder.forward                           lvdm/modules/networks/ae_modules.py:539-578
  ResnetBlock.forward (temb = None)         ae_modules.py:189-210
  AttnBlock.forward                         ae_modules.py:52-78
  Upsample.forward                          ae_modules.py:112-116
  Normalize / nonlinearity                  ae_modules.py:11-17  (GroupNorm(32, eps 1e-6), x * sigmoid(x))
"""
import torch
import torch.nn.functional as F


def _gn(x, sd, p):
    return F.group_norm(x, 32, sd[p + ".weight"], sd[p + ".bias"], 1e-6)


def _conv(x, sd, p, pad):
    return F.conv2d(x, sd[p + ".weight"], sd[p + ".bias"], padding=pad)


def _res(x, sd, p):
    h = _conv(F.silu(_gn(x, sd, p + ".norm1")), sd, p + ".conv1", 1)
    h = _conv(F.silu(_gn(h, sd, p + ".norm2")), sd, p + ".conv2", 1)
    if p + ".nin_shortcut.weight" in sd:
        x = _conv(x, sd, p + ".nin_shortcut", 0)
    return x + h


def _attn(x, sd, p):
    h = _gn(x, sd, p + ".norm")
    q, k, v = (_conv(h, sd, p + "." + n, 0) for n in ("q", "k", "v"))
    b, c, hh, ww = q.shape
    q = q.reshape(b, c, hh * ww).permute(0, 2, 1)
    k = k.reshape(b, c, hh * ww)
    w = torch.softmax(torch.bmm(q, k) * (int(c) ** -0.5), dim=2)
    v = v.reshape(b, c, hh * ww)
    h = torch.bmm(v, w.permute(0, 2, 1)).reshape(b, c, hh, ww)
    return x + _conv(h, sd, p + ".proj_out", 0)


def vae_decode(sd, z, n_levels, num_res_blocks):
    """AutoencoderKL.decode: z [n, zc, h, w] -> [n, out_ch, h * 2^(L-1), w * 2^(L-1)]."""
    sd = {k: v.float() for k, v in sd.items()}
    h = _conv(z.float(), sd, "post_quant_conv", 0)
    h = _conv(h, sd, "decoder.conv_in", 1)
    h = _res(h, sd, "decoder.mid.block_1")
    h = _attn(h, sd, "decoder.mid.attn_1")
    h = _res(h, sd, "decoder.mid.block_2")
    for lvl in reversed(range(n_levels)):
        for j in range(num_res_blocks + 1):
            h = _res(h, sd, f"decoder.up.{lvl}.block.{j}")
        if lvl != 0:
            h = F.interpolate(h, scale_factor=2.0, mode="nearest")
            h = _conv(h, sd, f"decoder.up.{lvl}.upsample.conv", 1)
    h = F.silu(_gn(h, sd, "decoder.norm_out"))
    return _conv(h, sd, "decoder.conv_out", 1)


def decode_first_stage_2DAE(sd, z, scale_factor, n_levels, num_res_blocks):
    """z [b, zc, t, h, w] -> [b, out_ch, t, 8h, 8w], frame by frame as the reference does."""
    z = z.float() / scale_factor
    return torch.cat([vae_decode(sd, z[:, :, i], n_levels, num_res_blocks).unsqueeze(2) for i in range(z.shape[2])], dim=2)


def random_vae_state_dict(shapes, seed=0):
    """Seeded decode-path weights: convs ~ N(0, 1 / fan_in), norm gains near 1, biases small; same on every machine."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shp in shapes.items():
        if len(shp) == 4:
            fan = shp[1] * shp[2] * shp[3]
            sd[k] = torch.randn(shp, generator=g) * (fan ** -0.5)
        elif ".norm" in k and k.endswith(".weight"):
            sd[k] = 1.0 + 0.1 * torch.randn(shp, generator=g)
        else:
            sd[k] = 0.05 * torch.randn(shp, generator=g)
    return sd


def vae_encode_moments(sd, x, n_levels, num_res_blocks):
    """AutoencoderKL.encode up to the posterior's parameters (autoencoder.py:97-102; Encoder.forward ae_modules.py:430-463;
    Downsample.forward :93-100): x [n, 3, H, W] -> moments [n, 2 zc, H / 2^(L-1), W / 2^(L-1)]."""
    sd = {k: v.float() for k, v in sd.items()}
    h = _conv(x.float(), sd, "encoder.conv_in", 1)
    for lvl in range(n_levels):
        for j in range(num_res_blocks):
            h = _res(h, sd, f"encoder.down.{lvl}.block.{j}")
        if lvl != n_levels - 1:
            h = F.pad(h, (0, 1, 0, 1), mode="constant", value=0)
            h = F.conv2d(h, sd[f"encoder.down.{lvl}.downsample.conv.weight"], sd[f"encoder.down.{lvl}.downsample.conv.bias"], stride=2)
    h = _res(h, sd, "encoder.mid.block_1")
    h = _attn(h, sd, "encoder.mid.attn_1")
    h = _res(h, sd, "encoder.mid.block_2")
    h = _conv(F.silu(_gn(h, sd, "encoder.norm_out")), sd, "encoder.conv_out", 1)
    return _conv(h, sd, "quant_conv", 0)


def posterior_sample(moments, noise):
    """DiagonalGaussianDistribution.sample (lvdm/distributions.py:24-43)."""
    mean, logvar = torch.chunk(moments, 2, dim=1)
    return mean + torch.exp(0.5 * torch.clamp(logvar, -30.0, 20.0)) * noise


def encode_first_stage_2DAE(sd, x, scale_factor, n_levels, num_res_blocks, noise):
    """ddpm3d.py:485-490 with get_first_stage_encoding (:458-465): x [b, 3, t, H, W], noise [b, zc, t, h, w]."""
    outs = [scale_factor * posterior_sample(vae_encode_moments(sd, x[:, :, i], n_levels, num_res_blocks), noise[:, :, i]).unsqueeze(2)
            for i in range(x.shape[2])]
    return torch.cat(outs, dim=2)

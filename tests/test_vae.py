"""First-stage decoder: oracle pinned to the reference fixture, the loader's parameter table, and GPU parity of
``AutoencoderKLDecoder`` (-> moca_vae_decode) with the fixture, the oracle and the full-size decode.

Bar: rel-L2 <= 2e-2 of the 16-bit-operand path against fp32 (north_star's tolerance for the tensor-core path).
"""
import hashlib
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

TOL = 2e-2
FULL = dict(ch=128, ch_mult=(1, 2, 4, 4), num_res_blocks=2, z_channels=4, out_ch=3)


def rel(a, b):
    a = a.float(); b = b.float()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _small():
    from moca_video_b200.vae import decoder_param_shapes
    from oracle.make_golden import VAE_SMALL
    from oracle.vae_ref import random_vae_state_dict
    return VAE_SMALL, random_vae_state_dict(decoder_param_shapes(**VAE_SMALL), seed=3)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_vae_restatement_matches_reference_fixture(tag):
    from oracle.vae_ref import vae_decode
    dd, sd = _small()
    g = np.load(os.path.join(GOLDEN, "vae_small.npz"))
    y = vae_decode(sd, torch.from_numpy(g[tag + "_z"]), len(dd["ch_mult"]), dd["num_res_blocks"])
    assert rel(y, torch.from_numpy(g[tag + "_y"])) < 1e-5


def test_vae_loader_expects_reference_state_dict():
    """moca_vae_create(ctx=NULL) lists what the packer consumes: the reference's decode-path keys at the shipped ddconfig
    (listing hashed from the reference's own Decoder in oracle/make_golden.py)."""
    from moca_video_b200.vae import AutoencoderKLDecoder, decoder_param_shapes
    g = np.load(os.path.join(GOLDEN, "vae_small.npz"))
    shapes = decoder_param_shapes(**FULL)
    listing = "\n".join(f"{k} {tuple(v)}" for k, v in sorted(shapes.items()))
    assert hashlib.sha256(listing.encode()).hexdigest() == str(g["full_shapes_sha256"])
    v = AutoencoderKLDecoder(FULL, _listing_only=True)
    assert v.expected_params() == {k: int(np.prod(s)) for k, s in shapes.items()}
    assert sum(v.expected_params().values()) == int(g["full_n_params"]) == 49490199
    dd, _ = _small()
    s = AutoencoderKLDecoder(dd, _listing_only=True)
    assert s.expected_params() == {k: int(np.prod(x)) for k, x in decoder_param_shapes(**dd).items()}


def _small_enc():
    from moca_video_b200.vae import encoder_param_shapes
    from oracle.make_golden import VAE_SMALL as dd
    from oracle.vae_ref import random_vae_state_dict
    return dd, random_vae_state_dict(encoder_param_shapes(dd["ch"], dd["ch_mult"], dd["num_res_blocks"], dd["z_channels"]), seed=4)


@pytest.mark.parametrize("tag", ["ea", "eb"])
def test_vae_encoder_restatement_matches_reference_fixture(tag):
    from oracle.vae_ref import vae_encode_moments
    dd, sd = _small_enc()
    g = np.load(os.path.join(GOLDEN, "vae_small.npz"))
    m = vae_encode_moments(sd, torch.from_numpy(g[tag + "_x"]), len(dd["ch_mult"]), dd["num_res_blocks"])
    assert rel(m, torch.from_numpy(g[tag + "_moments"])) < 1e-5


def test_vae_encoder_loader_expects_reference_state_dict():
    from moca_video_b200.vae import AutoencoderKLEncoder, encoder_param_shapes
    g = np.load(os.path.join(GOLDEN, "vae_small.npz"))
    shapes = encoder_param_shapes(128, (1, 2, 4, 4), 2, 4)
    listing = "\n".join(f"{k} {tuple(v)}" for k, v in sorted(shapes.items()))
    assert hashlib.sha256(listing.encode()).hexdigest() == str(g["full_enc_shapes_sha256"])
    e = AutoencoderKLEncoder(dict(FULL, in_channels=3, double_z=True), _listing_only=True)
    assert e.expected_params() == {k: int(np.prod(s)) for k, s in shapes.items()}
    assert sum(e.expected_params().values()) == int(g["full_enc_n_params"]) == 34163664


def test_vae_posterior_matches_reference_distribution():
    """DiagonalGaussian (vae.py) vs the reference's DiagonalGaussianDistribution on the same moments and noise."""
    from moca_video_b200.vae import DiagonalGaussian
    from oracle.vae_ref import posterior_sample
    g = torch.Generator().manual_seed(2)
    mom = torch.randn(2, 8, 4, 4, generator=g) * 20          # reaches both clamp bounds of logvar
    noise = torch.randn(2, 4, 4, 4, generator=g)
    post = DiagonalGaussian(mom)
    assert torch.equal(post.sample(noise), posterior_sample(mom, noise))
    assert torch.equal(post.mode(), mom[:, :4])
    torch.manual_seed(7)
    a = post.sample()
    torch.manual_seed(7)
    assert torch.equal(a, posterior_sample(mom, torch.randn(2, 4, 4, 4)))      # host generator, like the reference


def test_vae_unsupported_configs_fail_loudly():
    from moca_video_b200 import MocaError
    from moca_video_b200.vae import AutoencoderKLDecoder
    with pytest.raises(MocaError):
        AutoencoderKLDecoder(dict(FULL, tanh_out=True), _listing_only=True)
    with pytest.raises(MocaError):
        AutoencoderKLDecoder(dict(FULL, attn_resolutions=[16]), _listing_only=True)
    with pytest.raises(MocaError):
        AutoencoderKLDecoder(dict(FULL, ch=96), _listing_only=True)          # not a multiple of one GEMM k-block
    with pytest.raises(MocaError):
        AutoencoderKLDecoder(FULL, embed_dim=8, _listing_only=True)


# ---- GPU ---------------------------------------------------------------------------------------------------------------
@pytest.fixture
def _no_tf32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["a", "b"])
def test_vae_small_vs_reference_fixture(tag):
    from moca_video_b200.vae import AutoencoderKLDecoder
    dd, sd = _small()
    g = np.load(os.path.join(GOLDEN, "vae_small.npz"))
    dec = AutoencoderKLDecoder(dd, device="cuda").load_state_dict(sd)
    y = dec.decode(torch.from_numpy(g[tag + "_z"]).cuda())
    assert y.shape == g[tag + "_y"].shape
    assert rel(y.cpu(), torch.from_numpy(g[tag + "_y"])) < TOL


@pytest.mark.gpu
def test_vae_small_video_vs_oracle(_no_tf32):
    """decode_first_stage_2DAE: [b, c, t, h, w] latents, scale factor, frames grouped 3 per call vs one per call."""
    from moca_video_b200.vae import AutoencoderKLDecoder
    from oracle.vae_ref import decode_first_stage_2DAE
    dd, sd = _small()
    sd = {k: v.cuda() for k, v in sd.items()}
    gen = torch.Generator(device="cuda").manual_seed(9)
    z = torch.randn(2, 4, 5, 8, 16, device="cuda", generator=gen) * 0.18215
    ref = decode_first_stage_2DAE(sd, z, 0.18215, len(dd["ch_mult"]), dd["num_res_blocks"])
    dec = AutoencoderKLDecoder(dd, device="cuda", frames_per_call=3).load_state_dict(sd)
    y = dec.decode_first_stage_2DAE(z)
    assert y.shape == ref.shape == (2, 3, 5, 32, 64)
    assert rel(y, ref) < TOL
    one = AutoencoderKLDecoder(dd, device="cuda", frames_per_call=1).load_state_dict(sd)
    assert rel(one.decode_first_stage_2DAE(z), y) < 4e-3           # frames are independent; only rounding noise may differ


@pytest.mark.gpu
def test_vae_full_size_vs_oracle(_no_tf32):
    """The shipped ddconfig at the 320 x 512 output of the headline config, two latent frames."""
    from moca_video_b200.vae import AutoencoderKLDecoder, decoder_param_shapes
    from oracle.vae_ref import random_vae_state_dict, vae_decode
    sd = {k: v.cuda() for k, v in random_vae_state_dict(decoder_param_shapes(**FULL), seed=5).items()}
    gen = torch.Generator(device="cuda").manual_seed(11)
    z = torch.randn(2, 4, 40, 64, device="cuda", generator=gen)
    dec = AutoencoderKLDecoder(FULL, device="cuda", frames_per_call=2).load_state_dict(sd)
    y = dec.decode(z)
    assert y.shape == (2, 3, 320, 512)
    ref = torch.cat([vae_decode(sd, z[i:i + 1], 4, 2) for i in range(2)])
    assert torch.isfinite(y).all()
    assert rel(y, ref) < TOL
    assert dec.launches() > 0 and dec.flops() > 1e12


@pytest.mark.gpu
def test_vae_rejects_bad_input():
    from moca_video_b200 import MocaError
    from moca_video_b200.vae import AutoencoderKLDecoder
    dd, sd = _small()
    dec = AutoencoderKLDecoder(dd, device="cuda")
    with pytest.raises(MocaError):
        dec.decode(torch.zeros(1, 4, 8, 8, device="cuda"))           # weights not loaded
    dec.load_state_dict(sd)
    with pytest.raises(MocaError):
        dec.decode(torch.zeros(1, 3, 8, 8, device="cuda"))           # wrong channel count
    with pytest.raises(MocaError):
        dec.decode(torch.zeros(1, 4, 6, 6, device="cuda"))           # H * W not a multiple of 64 (attention k-blocks)


class _StandIn(torch.nn.Module):
    """A module tree with the attribute names attach_vae reads from the reference AutoencoderKL (decoder.ch, .num_res_blocks,
    .up[i].block / .attn / .upsample.conv, conv_in, conv_out, post_quant_conv), built from the parameter table; the GPU box has no
    reference tree to construct the real class from."""

    def __init__(self, dd, sd):
        super().__init__()
        nn = torch.nn

        mods = {}
        for k, v in sd.items():
            name, leaf = k.rsplit(".", 1)
            mods.setdefault(name, {})[leaf] = v
        # ModuleLists need their entries in index order
        for name in sorted(mods, key=lambda s: [int(p) if p.isdigit() else p for p in s.split(".")]):
            w = mods[name]["weight"]
            m = nn.Conv2d(w.shape[1], w.shape[0], w.shape[2], padding=w.shape[2] // 2) if w.dim() == 4 else nn.GroupNorm(32, w.shape[0], eps=1e-6)
            parts = name.split(".")
            root = self
            for i, p in enumerate(parts[:-1]):
                if p.isdigit():
                    while len(root) <= int(p):
                        root.append(nn.Module())
                    root = root[int(p)]
                else:
                    if not hasattr(root, p):
                        root.add_module(p, nn.ModuleList() if p in ("up", "block") else nn.Module())
                    root = getattr(root, p)
            root.add_module(parts[-1], m)
        for u in self.decoder.up:
            u.add_module("attn", nn.ModuleList())
        self.decoder.ch = dd["ch"]
        self.decoder.num_res_blocks = dd["num_res_blocks"]
        self.decoder.give_pre_end = False
        self.decoder.tanh_out = False
        self.quant_conv = nn.Conv2d(8, 8, 1)                       # a key outside the decode path: must be ignored by the loader
        self.load_state_dict(sd, strict=False)


@pytest.mark.gpu
def test_attach_vae_reads_the_reference_module_tree(_no_tf32):
    from moca_video_b200.patch import _ae_decode, _decode_first_stage_2DAE
    from moca_video_b200.vae import attach_vae
    from oracle.vae_ref import decode_first_stage_2DAE, vae_decode
    import types
    dd, sd = _small()
    ae = _StandIn(dd, sd).cuda()
    eng = attach_vae(ae)
    assert (eng.ch, eng.ch_mult, eng.num_res_blocks, eng.z_channels, eng.out_ch) == (64, (1, 2, 2), 1, 4, 3)
    assert attach_vae(ae) is eng
    sdc = {k: v.cuda() for k, v in sd.items()}
    gen = torch.Generator(device="cuda").manual_seed(3)
    z = torch.randn(2, 4, 8, 8, device="cuda", generator=gen)
    assert rel(_ae_decode(ae, z), vae_decode(sdc, z, 3, 1)) < TOL
    ld = types.SimpleNamespace(first_stage_model=ae, scale_factor=0.5)
    zz = torch.randn(1, 4, 3, 8, 8, device="cuda", generator=gen)
    assert rel(_decode_first_stage_2DAE(ld, zz), decode_first_stage_2DAE(sdc, zz, 0.5, 3, 1)) < TOL


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["ea", "eb"])
def test_vae_encoder_small_vs_reference_fixture(tag):
    from moca_video_b200.vae import AutoencoderKLEncoder
    dd, sd = _small_enc()
    g = np.load(os.path.join(GOLDEN, "vae_small.npz"))
    enc = AutoencoderKLEncoder(dict(dd, in_channels=3, double_z=True), device="cuda").load_state_dict(sd)
    m = enc.encode_moments(torch.from_numpy(g[tag + "_x"]).cuda())
    assert m.shape == g[tag + "_moments"].shape
    assert rel(m.cpu(), torch.from_numpy(g[tag + "_moments"])) < TOL


@pytest.mark.gpu
def test_vae_encode_first_stage_vs_oracle(_no_tf32):
    from moca_video_b200.vae import AutoencoderKLEncoder
    from oracle.vae_ref import encode_first_stage_2DAE
    dd, sd = _small_enc()
    sd = {k: v.cuda() for k, v in sd.items()}
    gen = torch.Generator(device="cuda").manual_seed(13)
    x = torch.rand(1, 3, 3, 32, 64, device="cuda", generator=gen) * 2 - 1
    noise = torch.randn(1, 4, 3, 8, 16, device="cuda", generator=gen)
    enc = AutoencoderKLEncoder(dict(dd, in_channels=3, double_z=True), device="cuda", scale_factor=0.18215).load_state_dict(sd)
    z = enc.encode_first_stage_2DAE(x, noise=noise)
    ref = encode_first_stage_2DAE(sd, x, 0.18215, 3, 1, noise)
    assert z.shape == ref.shape == (1, 4, 3, 8, 16)
    assert rel(z, ref) < TOL


@pytest.mark.gpu
def test_vae_encoder_full_size_vs_oracle(_no_tf32):
    """Shipped ddconfig, one 320 x 512 frame (the DAVIS branch of prepare_latents, funcs.py:38-47): every wide-row / stride-2
    tensor-map geometry of the encoder."""
    from moca_video_b200.vae import AutoencoderKLEncoder, encoder_param_shapes
    from oracle.vae_ref import random_vae_state_dict, vae_encode_moments
    sd = {k: v.cuda() for k, v in random_vae_state_dict(encoder_param_shapes(128, (1, 2, 4, 4), 2, 4), seed=6).items()}
    gen = torch.Generator(device="cuda").manual_seed(17)
    x = torch.rand(1, 3, 320, 512, device="cuda", generator=gen) * 2 - 1
    enc = AutoencoderKLEncoder(dict(FULL, in_channels=3, double_z=True), device="cuda").load_state_dict(sd)
    m = enc.encode_moments(x)
    assert m.shape == (1, 8, 40, 64)
    assert rel(m, vae_encode_moments(sd, x, 4, 2)) < TOL


@pytest.mark.gpu
def test_davis_branches_of_prepare_and_shift(_no_tf32):
    """prepare_latents(data=...) and shift_latents(davis_data=...) (funcs.py:38-47, 100-116): the queue starts from the encoded
    frames, the shift anchors on the encoded LAST frame, the mask queue moves with the latents."""
    import types
    from moca_video_b200 import fifo, ops
    from moca_video_b200.vae import AutoencoderKLEncoder
    from oracle.vae_ref import encode_first_stage_2DAE
    dd, sd = _small_enc()
    enc = AutoencoderKLEncoder(dict(dd, in_channels=3, double_z=True), device="cuda", scale_factor=0.18215).load_state_dict(sd)
    model = types.SimpleNamespace(encode_first_stage_2DAE=enc.encode_first_stage_2DAE)
    gen = torch.Generator(device="cuda").manual_seed(21)
    frames = torch.rand(1, 4, 4, 64, 64, device="cuda", generator=gen) * 2 - 1             # RGBA: alpha is dropped
    masks = (torch.rand(1, 1, 12, 16, 16, device="cuda", generator=gen) > 0.5).float()
    fa = types.SimpleNamespace(num_inference_steps=8, lookahead_denoising=True, video_length=8)
    sampler = types.SimpleNamespace(ddim_alphas=np.linspace(0.9, 0.1, 8))
    qn = torch.randn(1, 4, 12, 16, 16, device="cuda", generator=gen)
    torch.manual_seed(5)                                                                    # the posterior's host-side draw
    q = fifo.prepare_latents(fa, None, sampler, model, data=(frames, masks), noise=qn)
    torch.manual_seed(5)
    z = enc.encode_first_stage_2DAE(frames[:, :3])
    assert torch.equal(q, fifo.prepare_latents(fa, None, sampler, initial_latents=z, noise=qn))
    sdc = {k: v.cuda() for k, v in sd.items()}
    torch.manual_seed(5)
    host_noise = torch.cat([torch.randn(1, 4, 16, 16).unsqueeze(2) for _ in range(4)], dim=2).cuda()
    assert rel(z, encode_first_stage_2DAE(sdc, frames[:, :3], 0.18215, 3, 1, host_noise)) < TOL
    # shift: same tail as the latent-anchored shift would give with the encoded last frame in slot 0
    nn_ = torch.randn(1, 4, 1, 16, 16, device="cuda", generator=gen)
    torch.manual_seed(6)
    out, (f2, m2) = fifo.shift_latents(q.clone(), (frames, masks), model, new_noise=nn_)
    torch.manual_seed(6)
    anchor = enc.encode_first_stage_2DAE(frames[:, :3, -1:])
    fake = q.clone()
    fake[:, :, 0] = anchor[:, :, 0]
    want = fifo.shift_latents(fake, new_noise=nn_)
    assert torch.equal(out[:, :, -1], want[:, :, -1]) and torch.equal(out[:, :, :-1], q[:, :, 1:])
    assert f2 is frames and torch.equal(m2[:, :, :-1], masks[:, :, 1:]) and torch.equal(m2[:, :, -1], masks[:, :, -1])

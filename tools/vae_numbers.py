"""Parity and timing numbers of the first-stage autoencoder at the shipped ddconfig / 320 x 512 (for DESIGN.md / profiles)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
from moca_video_b200.vae import AutoencoderKLDecoder, AutoencoderKLEncoder, decoder_param_shapes, encoder_param_shapes
from oracle.vae_ref import random_vae_state_dict, vae_decode, vae_encode_moments

FULL = dict(ch=128, ch_mult=(1, 2, 4, 4), num_res_blocks=2, z_channels=4, out_ch=3)
rel = lambda a, b: float((a.float() - b.float()).norm() / b.float().norm())
out = {}
sd = {k: v.cuda() for k, v in random_vae_state_dict(decoder_param_shapes(**FULL), seed=5).items()}
g = torch.Generator(device="cuda").manual_seed(11)
z = torch.randn(16, 4, 40, 64, device="cuda", generator=g)
ref = vae_decode(sd, z[:1], 4, 2)
def timeit(fn, n=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for fpc in (1, 2, 4):
    dec = AutoencoderKLDecoder(FULL, device="cuda", frames_per_call=fpc).load_state_dict(sd)
    y = dec.decode(z[:fpc])
    out[f"decode_rel_l2_fpc{fpc}"] = rel(y[:1], ref)
    out[f"decode_ms_per_frame_fpc{fpc}"] = timeit(lambda: dec.decode(z)) / 16
    out[f"decode_workspace_gib_fpc{fpc}"] = dec.workspace_bytes() / 2**30
    out["decode_tflop_per_frame"] = dec.flops() / fpc / 1e12
    dec.close()
# torch eager fp32 / bf16-autocast decode of the same frame for scale (same modules as the reference's, restated)
out["torch_fp32_decode_ms_per_frame"] = timeit(lambda: vae_decode(sd, z[:1], 4, 2))
esd = {k: v.cuda() for k, v in random_vae_state_dict(encoder_param_shapes(128, (1, 2, 4, 4), 2, 4), seed=6).items()}
x = torch.rand(4, 3, 320, 512, device="cuda", generator=g) * 2 - 1
enc = AutoencoderKLEncoder(dict(FULL, in_channels=3, double_z=True), device="cuda", frames_per_call=2).load_state_dict(esd)
m = enc.encode_moments(x[:2])
out["encode_rel_l2"] = rel(m[:1], vae_encode_moments(esd, x[:1], 4, 2))
out["encode_ms_per_frame"] = timeit(lambda: enc.encode_moments(x)) / 4
out["encode_tflop_per_frame"] = enc.flops() / 2 / 1e12
out["torch_fp32_encode_ms_per_frame"] = timeit(lambda: vae_encode_moments(esd, x[:1], 4, 2))
print(json.dumps(out, indent=1))
json.dump(out, open("gpurun_out/vae_numbers.json", "w"), indent=1)

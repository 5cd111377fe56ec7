"""Device timing of the first-stage decode at the headline size (16 frames of 40 x 64 latents -> 320 x 512) and a per-op table.
Usage: python tools/vae_profile.py [frames_per_call] [out.csv]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from moca_video_b200.vae import AutoencoderKLDecoder, decoder_param_shapes

FULL = dict(ch=128, ch_mult=(1, 2, 4, 4), num_res_blocks=2, z_channels=4, out_ch=3)
fpc = int(sys.argv[1]) if len(sys.argv) > 1 else 2
g = torch.Generator().manual_seed(5)
sd = {}
for k, shp in decoder_param_shapes(**FULL).items():
    if len(shp) == 4:
        sd[k] = torch.randn(shp, generator=g) * (shp[1] * shp[2] * shp[3]) ** -0.5
    elif ".norm" in k and k.endswith(".weight"):
        sd[k] = 1.0 + 0.1 * torch.randn(shp, generator=g)
    else:
        sd[k] = 0.05 * torch.randn(shp, generator=g)
dec = AutoencoderKLDecoder(FULL, device="cuda", frames_per_call=fpc).load_state_dict(sd)
z = torch.randn(1, 4, 16, 40, 64, device="cuda") * 0.18215
for _ in range(2):
    y = dec.decode_first_stage_2DAE(z)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    y = dec.decode_first_stage_2DAE(z)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
print(f"decode_first_stage_2DAE 16 frames 320x512, {fpc} frames/call: {ms:.2f} ms ({ms / 16:.2f} ms/frame), "
      f"{dec.flops() / fpc / 1e12:.3f} TFLOP/frame -> {dec.flops() / fpc * 16 / ms / 1e9:.1f} TFLOP/s, launches/call {dec.launches()}, "
      f"workspace {dec.workspace_bytes() / 2**30:.2f} GiB")
rows = dec.profile(3)
tot = sum(r[0] for r in rows)
print(f"per-op sum {tot / 1e3:.2f} ms per call")
for us, gf, label in sorted(rows, reverse=True)[:25]:
    print(f"{us:9.1f} us {gf / us * 1e3 if us else 0:8.1f} TF/s  {label}")
if len(sys.argv) > 2:
    with open(sys.argv[2], "w") as f:
        f.write("us,gflop,label\n")
        for us, gf, label in rows:
            f.write(f"{us:.2f},{gf:.3f},{label}\n")

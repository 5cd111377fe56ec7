"""Per-op device timing of the full-size UNet forward (moca_unet_profile): python tools/profile_forward.py [B] [out.csv]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moca_video_b200.arch import UNetConfig, random_state_dict
from moca_video_b200.unet import UNet3D

dev = "cuda"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 2
out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/profile_ops.csv"
cfg = UNetConfig.videocrafter2()
net = UNet3D(device=dev).load_state_dict(random_state_dict(cfg, seed=0, device=dev))
g = torch.Generator(device=dev).manual_seed(1)
x = torch.randn(B, 4, 16, 40, 64, device=dev, generator=g)
c = torch.randn(B, 77, 1024, device=dev, generator=g)
ts = list(range(0, 16 * 60, 60)) * B
for _ in range(3):
    y = net(x, ts, c, fps=10)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    net(x, ts, c, fps=10)
e1.record(); torch.cuda.synchronize()
print(f"B={B} forward {e0.elapsed_time(e1) / 10:.3f} ms (graph)")
rows = net.profile(5)
tot = sum(r[0] for r in rows)
with open(out, "w") as f:
    f.write("us,gflop,tflops,label\n")
    for us, gf, label in rows:
        f.write(f"{us:.2f},{gf:.2f},{(gf / us / 1e3 if us > 0 else 0):.1f},{label}\n")
print(f"sum of per-op times {tot / 1e3:.3f} ms over {len(rows)} ops")
import collections, re
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for us, gf, label in rows:
    kind = label.split(" ")[1] if " " in label else label
    a = agg[kind]; a[0] += 1; a[1] += us; a[2] += gf
for k, (n, us, gf) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k:12s} n={n:4d} {us / 1e3:8.3f} ms {100 * us / tot:5.1f}%  {gf / us / 1e3 if us else 0:7.1f} TFLOP/s")

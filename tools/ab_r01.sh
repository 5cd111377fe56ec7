#!/bin/bash
# same-box comparison of the round-1 package (gpurun_tmp/r01_pkg, built from commit 379ab4a) with the working tree
mkdir -p gpurun_out
cat > /tmp/prof_r01.py <<'PY'
import os, sys, torch
sys.path.insert(0, sys.argv[1])
out = sys.argv[2]
from moca_video_b200.arch import UNetConfig, random_state_dict
from moca_video_b200.unet import UNet3D
dev = "cuda"
cfg = UNetConfig.videocrafter2()
net = UNet3D(device=dev).load_state_dict(random_state_dict(cfg, seed=0, device=dev))
g = torch.Generator(device=dev).manual_seed(1)
x = torch.randn(2, 4, 16, 40, 64, device=dev, generator=g)
c = torch.randn(2, 77, 1024, device=dev, generator=g)
ts = list(range(0, 16 * 60, 60)) * 2
for _ in range(3):
    y = net(x, ts, c, fps=10)
torch.cuda.synchronize()
best = 1e9
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        net(x, ts, c, fps=10)
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 20)
print(f"forward {best:.3f} ms (graph, best of 3 x 20)")
rows = net.profile(5)
with open(out, "w") as f:
    f.write("us,gflop,tflops,label\n")
    for us, gf, label in rows:
        f.write(f"{us:.2f},{gf:.2f},0,{label}\n")
import collections
agg = collections.defaultdict(float)
for us, gf, label in rows:
    agg[label.split(" ")[1]] += us
print("  " + "  ".join(f"{k}={v / 1e3:.2f}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
PY
mkdir -p /tmp/r01root && rm -rf /tmp/r01root/moca_video_b200 && cp -r gpurun_tmp/r01_pkg /tmp/r01root/moca_video_b200
for i in 0 1; do
  echo "== round 1"; python /tmp/prof_r01.py /tmp/r01root gpurun_out/ops_r01.csv
  echo "== now";     python /tmp/prof_r01.py . gpurun_out/ops_now.csv
done

"""One full-size b=2 UNet forward (for ncu captures): python tools/one_forward.py [n_forwards]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moca_video_b200.arch import UNetConfig, random_state_dict
from moca_video_b200.unet import UNet3D
dev = "cuda"
cfg = UNetConfig.videocrafter2()
net = UNet3D(device=dev).load_state_dict(random_state_dict(cfg, seed=0, device=dev))
g = torch.Generator(device=dev).manual_seed(1)
x = torch.randn(2, 4, 16, 40, 64, device=dev, generator=g)
c = torch.randn(2, 77, 1024, device=dev, generator=g)
ts = list(range(0, 16 * 60, 60)) * 2
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 1):
    y = net(x, ts, c, fps=10)
torch.cuda.synchronize()
print("ok", float(y.abs().mean()))

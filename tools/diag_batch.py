"""Why does a b=2 forward differ from two b=1 forwards?  Determinism + per-tap comparison."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moca_video_b200.arch import UNetConfig, random_state_dict
from moca_video_b200.unet import UNet3D

dev = "cuda"
rel = lambda a, b: float((a.float() - b.float()).norm() / (b.float().norm() + 1e-30))
full = len(sys.argv) > 1 and sys.argv[1] == "full"
cfg = UNetConfig.videocrafter2() if full else UNetConfig(model_channels=64, context_dim=64)
sd = random_state_dict(cfg, seed=0, device=dev)
net = UNet3D(model_channels=cfg.model_channels, context_dim=cfg.context_dim, device=dev).load_state_dict(sd)
g = torch.Generator(device=dev).manual_seed(9)
H, W, L = (40, 64, 77) if full else (8, 16, 11)
x = torch.randn(1, 4, 16, H, W, device=dev, generator=g)
c = torch.randn(1, L, cfg.context_dim, device=dev, generator=g)
ts = list(range(100, 900, 50))
y1 = net(x, ts, c, fps=10); y2 = net(x, ts, c, fps=10)
print("b=1 run-to-run rel", rel(y1, y2), "equal", torch.equal(y1, y2))
e2 = net(torch.cat([x, x]), ts + ts, torch.cat([c, c]), fps=10)
print("b=2 sample0 vs sample1", rel(e2[0], e2[1]), " sample0 vs b=1", rel(e2[0:1], y1), " sample1 vs b=1", rel(e2[1:2], y1))
net.set_taps(True)
net(x, ts, c, fps=10); t1 = net.taps()
net(torch.cat([x, x]), ts + ts, torch.cat([c, c]), fps=10); t2 = net.taps()
for k in t1:
    n = t1[k].shape[0]
    print(f"  {k:30s} s0 {rel(t2[k][:n], t1[k]):.3e}  s1 {rel(t2[k][n:], t1[k]):.3e}")

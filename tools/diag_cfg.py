"""Localise the difference between the b=2 CFG launch (ctx_len 154/77) and separate b=1 forwards, per layer tap."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moca_video_b200.arch import UNetConfig, random_state_dict
from moca_video_b200.unet import UNet3D

dev = "cuda"
rel = lambda a, b: float((a.float() - b.float()).norm() / (b.float().norm() + 1e-30))
cfg = UNetConfig.videocrafter2()
sd = random_state_dict(cfg, seed=0, device=dev)
net = UNet3D(device=dev).load_state_dict(sd)
g = torch.Generator(device=dev).manual_seed(9)
x = torch.randn(1, 4, 16, 40, 64, device=dev, generator=g)
c = torch.randn(1, 154, 1024, device=dev, generator=g)
u = torch.randn(1, 77, 1024, device=dev, generator=g)
ts = list(range(100, 900, 50))
poison = float(sys.argv[1]) if len(sys.argv) > 1 else 1e4
ctx = torch.zeros(2, 154, 1024, device=dev)
ctx[0] = c[0]; ctx[1, :77] = u[0]; ctx[1, 77:] = poison
e_u = net(x, ts, u, fps=10); e_u2 = net(x, ts, u, fps=10)
print("b=1 L77 run-to-run", rel(e_u, e_u2))
e_up = net(x, ts, ctx[1:2], fps=10, ctx_len=[77])
print("b=1 Lmax154 len77 poisoned vs L77", rel(e_up, e_u))
e2 = net(torch.cat([x, x]), ts + ts, ctx, fps=10, ctx_len=[154, 77])
print("b=2 sample1 vs L77", rel(e2[1:2], e_u), " vs padded b=1", rel(e2[1:2], e_up))
net.set_taps(True)
net(x, ts, u, fps=10); t1 = net.taps()
net(x, ts, ctx[1:2], fps=10, ctx_len=[77]); t3 = net.taps()
net(torch.cat([x, x]), ts + ts, ctx, fps=10, ctx_len=[154, 77]); t2 = net.taps()
for k in t1:
    n = t1[k].shape[0]
    print(f"  {k:30s} batch-s1 vs L77 {rel(t2[k][n:], t1[k]):.3e}   padded-b1 vs L77 {rel(t3[k], t1[k]):.3e}")

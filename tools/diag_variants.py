"""Per-tap comparison of two builds / environment variants of the small UNet against the fp32 oracle.
   python tools/diag_variants.py A=MOCA_GEMM_TEPI=1 B=MOCA_GEMM_TEPI=0"""
import os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import sys, torch, numpy as np
sys.path.insert(0, %r)
from moca_video_b200.arch import UNetConfig, random_state_dict
from moca_video_b200.unet import UNet3D
from oracle.unet_ref import unet_forward
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
cfg = UNetConfig(model_channels=64, context_dim=64)
sd = random_state_dict(cfg, seed=0, device="cuda")
net = UNet3D(model_channels=64, context_dim=64, device="cuda").load_state_dict(sd)
g = torch.Generator(device="cuda").manual_seed(3)
x = torch.randn(1, 4, 16, 32, 32, device="cuda", generator=g)
c = torch.randn(1, 77, 64, device="cuda", generator=g)
ts = list(range(0, 960, 60))
rt = {}
y_ref = unet_forward(sd, cfg, x, torch.tensor(ts, device="cuda"), c, fps=10, taps=rt)
net.set_taps(True)
y = net(x, ts, c, fps=10)
taps = net.taps()
out = {"y": y.cpu().numpy(), "y_ref": y_ref.cpu().numpy()}
names = []
for k, v in taps.items():
    out["tap_" + k] = v.cpu().numpy(); out["ref_" + k] = rt[k].cpu().numpy(); names.append(k)
out["names"] = np.array(names)
np.savez(sys.argv[1], **out)
"""
def run(tag, envs):
    env = dict(os.environ)
    for kv in envs.split(","):
        if kv:
            k, v = kv.split("=", 1); env[k] = v
    path = f"/tmp/diag_{tag}.npz"
    r = subprocess.run([sys.executable, "-c", CHILD % ROOT, path], env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return np.load(path)
rel = lambda a, b: float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))
specs = dict(a.split("=", 1) for a in sys.argv[1:])
A, B = run("A", specs.get("A", "")), run("B", specs.get("B", ""))
print(f"{'tap':32s} {'A vs ref':>10s} {'B vs ref':>10s} {'A vs B':>10s}")
for k in A["names"]:
    print(f"{k:32s} {rel(A['tap_'+k], A['ref_'+k]):10.3e} {rel(B['tap_'+k], B['ref_'+k]):10.3e} {rel(A['tap_'+k], B['tap_'+k]):10.3e}")
print(f"{'y':32s} {rel(A['y'], A['y_ref']):10.3e} {rel(B['y'], B['y_ref']):10.3e} {rel(A['y'], B['y']):10.3e}")

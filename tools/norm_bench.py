"""Device timing of the LayerNorm / GroupNorm primitives at the UNet's shapes: python tools/norm_bench.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moca_video_b200 import ops

dev = "cuda"

def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

for rows, C in [(81920, 320), (20480, 640), (5120, 1280)]:
    x = torch.randn(rows, C, device=dev)
    ga = torch.randn(C, device=dev); be = torch.randn(C, device=dev)
    us = timeit(lambda: ops.layernorm_op16(x, ga, be))
    print(f"layernorm rows={rows} C={C}: {us:7.1f} us  {rows * C * 6 / us / 1e3:7.0f} GB/s (incl. output allocation)")
    us = timeit(lambda: ops.groupnorm_op16(x, ga, be, rows // 32, 1e-5, 1))
    print(f"groupnorm (stats + apply) rows={rows} C={C}: {us:7.1f} us  {rows * C * 10 / us / 1e3:7.0f} GB/s")

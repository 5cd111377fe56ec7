"""Device timing of the attention primitive at the UNet's spatial shapes (b=2 launch): python tools/attn_bench.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moca_video_b200 import ops

dev = "cuda"
for (batch, heads, Lq, Lk, Lv) in [(32, 5, 2560, 2560, 2560), (32, 10, 640, 640, 640), (32, 20, 160, 160, 160), (32, 5, 2560, 77, 77), (32, 10, 640, 77, 77)]:
    g = torch.Generator(device=dev).manual_seed(0)
    C = heads * 64
    qkv = torch.randn(batch * Lq, 3 * C, device=dev, generator=g).to(ops.op_dtype())
    if Lk == Lq:
        q, k, v = qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:]
    else:
        q = qkv[:, :C]
        kv = torch.randn(batch * Lk, 2 * C, device=dev, generator=g).to(ops.op_dtype())
        k, v = kv[:, :C], kv[:, C:]
    for _ in range(3):
        ops.attention_op16(q, k, v, batch, heads, Lq, Lk, Lv)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        ops.attention_op16(q, k, v, batch, heads, Lq, Lk, Lv)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    fl = 4.0 * batch * heads * Lq * Lv * 64
    print(f"attn batch={batch} heads={heads} Lq={Lq} Lk={Lv}: {us:8.1f} us  {fl / us / 1e6:7.1f} TFLOP/s")

"""Parity and timing of the ViT-H-14 text tower (layer = penultimate) for a CFG prompt pair."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
torch.backends.cuda.matmul.allow_tf32 = False
from moca_video_b200.text import OpenCLIPTextEncoder, text_param_shapes
from oracle.text_ref import random_text_state_dict, text_encode
sd = {k: v.cuda() for k, v in random_text_state_dict(text_param_shapes(49408, 77, 1024, 16, 23), seed=9).items()}
tokens = torch.randint(1, 49405, (2, 77), generator=torch.Generator().manual_seed(3))
enc = OpenCLIPTextEncoder(device="cuda", layer="penultimate").load_state_dict(sd)
def timeit(fn, n=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
tc = tokens.cuda()
out = {"rel_l2": float((enc.encode_with_transformer(tokens) - text_encode(sd, tc, 16, 23)).norm() / text_encode(sd, tc, 16, 23).norm()),
       "ms_b2": timeit(lambda: enc.encode_with_transformer(tc)), "torch_fp32_ms_b2": timeit(lambda: text_encode(sd, tc, 16, 23)),
       "launches": enc.launches(), "weight_mb": sum(v.numel() for k, v in sd.items() if "embedding" not in k) * 2 / 1e6}
rows = enc.profile(5)
out["per_op_sum_ms"] = sum(r[0] for r in rows) / 1e3
print(json.dumps(out, indent=1))
for us, gf, label in sorted(rows, reverse=True)[:8]:
    print(f"{us:8.1f} us  {label}")
json.dump(out, open("gpurun_out/text_numbers.json", "w"), indent=1)

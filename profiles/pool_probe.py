"""Five fused pool + normalise + append launches, for ncu: python profiles/pool_probe.py [bf16|f32] [full|ragged]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import docqa_b200 as bfk  # noqa: E402

dt = torch.float32 if (len(sys.argv) > 1 and sys.argv[1] == "f32") else torch.bfloat16
ragged = len(sys.argv) > 2 and sys.argv[2] == "ragged"
B, T, H = 16384, 256, 384
g = torch.Generator(device="cuda").manual_seed(1234)
tok = torch.randn((B, T, H), generator=g, device="cuda").to(dt)
if ragged:
    ln = torch.randint(16, T + 1, (B,), generator=g, device="cuda")
    mask = (torch.arange(T, device="cuda")[None, :] < ln[:, None]).to(torch.int64)
else:
    mask = torch.ones((B, T), dtype=torch.int64, device="cuda")
idx = bfk.IndexFlatL2(H)
idx.reserve(B * 6)
for i in range(5):
    idx.add_pooled(tok, mask)
torch.cuda.synchronize()
print("ok", idx.ntotal, "valid tokens per chunk", float(mask.sum()) / B)

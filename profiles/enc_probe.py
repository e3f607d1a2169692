"""Encoder pipeline probe: python profiles/enc_probe.py [B] [T] -- two encode_add calls (for ncu launch lists)."""
import sys

import torch

sys.path.insert(0, "/root/repo")
import docqa_b200 as bfk  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
T = int(sys.argv[2]) if len(sys.argv) > 2 else 256
enc = bfk.MiniLMEncoder(device=0)
g = torch.Generator().manual_seed(0)
H, I = 384, 1536
sd = {"embeddings.word_embeddings.weight": 0.05 * torch.randn((30522, H), generator=g),
      "embeddings.position_embeddings.weight": 0.05 * torch.randn((512, H), generator=g),
      "embeddings.token_type_embeddings.weight": 0.05 * torch.randn((2, H), generator=g),
      "embeddings.LayerNorm.weight": torch.ones(H), "embeddings.LayerNorm.bias": torch.zeros(H)}
for li in range(6):
    p = f"encoder.layer.{li}."
    for name, shape in (("attention.self.query", (H, H)), ("attention.self.key", (H, H)), ("attention.self.value", (H, H)),
                        ("attention.output.dense", (H, H)), ("intermediate.dense", (I, H)), ("output.dense", (H, I))):
        sd[p + name + ".weight"] = 0.05 * torch.randn(shape, generator=g)
        sd[p + name + ".bias"] = 0.05 * torch.randn(shape[0], generator=g)
    for name in ("attention.output.LayerNorm", "output.LayerNorm"):
        sd[p + name + ".weight"] = torch.ones(H)
        sd[p + name + ".bias"] = torch.zeros(H)
enc.load_state_dict(sd)
ids = torch.randint(0, 30522, (B, T), device="cuda")
mask = torch.ones((B, T), dtype=torch.int64, device="cuda")
idx = bfk.IndexFlatL2(H)
idx.reserve(4 * B)
for _ in range(2):
    enc.encode_add(idx, ids, mask)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
enc.encode_add(idx, ids, mask)
e1.record()
torch.cuda.synchronize()
print("B", B, "T", T, "ms", e0.elapsed_time(e1), "chunks/s", B / e0.elapsed_time(e1) * 1e3)

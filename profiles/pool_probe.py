import sys, torch
sys.path.insert(0, '/root/repo')
import docqa_b200 as bfk
B, T, H = 16384, 256, 384
tok = torch.randn((B, T, H), device="cuda").to(torch.bfloat16)
mask = torch.ones((B, T), dtype=torch.int64, device="cuda")
idx = bfk.IndexFlatL2(H); idx.reserve(B * 6)
for i in range(5):
    idx.add_pooled(tok, mask)
torch.cuda.synchronize()
print("ok", idx.ntotal)

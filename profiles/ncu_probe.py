"""One tensor-path search at a chosen batch size, for `ncu --set full -k regex:scan_tc_kernel -s 1 -c 1`.

    python profiles/ncu_probe.py <nq> [rows] [d] [k]
"""
import sys

import torch

sys.path.insert(0, "/root/repo")
import docqa_b200 as bfk  # noqa: E402

nq = int(sys.argv[1])
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
d = int(sys.argv[3]) if len(sys.argv) > 3 else 384
k = int(sys.argv[4]) if len(sys.argv) > 4 else 10
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1234)
idx = bfk.IndexFlatL2(d, device=0)
idx.reserve(n)
for r0 in range(0, n, 1 << 20):
    x = torch.randn((min(1 << 20, n - r0), d), generator=g, device=dev)
    x /= x.norm(dim=1, keepdim=True)
    idx.add(x)
q = torch.randn((nq, d), generator=g, device=dev)
q /= q.norm(dim=1, keepdim=True)
st = bfk.SearchStats()
for _ in range(3):
    D, I = idx.search(q, k, algo="tensor", stats=st)
torch.cuda.synchronize()
print("nq", nq, "scan_ms", st.scan_ms, "fallback", st.n_fallback)

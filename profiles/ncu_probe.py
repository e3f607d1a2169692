"""A few searches of one shape, for `ncu ... -k regex:<kernel> -s <skip> -c 1 python profiles/ncu_probe.py ...`

    python profiles/ncu_probe.py <nq> [rows] [d] [k] [metric: l2|ip] [storage: f32|bf16] [algo: tensor|exact|auto]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import docqa_b200 as bfk  # noqa: E402

nq = int(sys.argv[1])
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
d = int(sys.argv[3]) if len(sys.argv) > 3 else 384
k = int(sys.argv[4]) if len(sys.argv) > 4 else 10
metric = bfk.METRIC_INNER_PRODUCT if (len(sys.argv) > 5 and sys.argv[5] == "ip") else bfk.METRIC_L2
storage = sys.argv[6] if len(sys.argv) > 6 else "f32"
algo = sys.argv[7] if len(sys.argv) > 7 else "tensor"
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1234)
idx = bfk.IndexFlat(d, metric, device=0, storage=storage)
idx.reserve(n)
for r0 in range(0, n, 1 << 20):
    x = torch.randn((min(1 << 20, n - r0), d), generator=g, device=dev)
    x /= x.norm(dim=1, keepdim=True)
    idx.add(x)
    del x
q = torch.randn((nq, d), generator=g, device=dev)
q /= q.norm(dim=1, keepdim=True)
st = bfk.SearchStats()
for _ in range(3):
    D, I = idx.search(q, k, algo=algo, stats=st)
torch.cuda.synchronize()
print("nq", nq, "rows", n, "d", d, "k", k, "scan_ms", st.scan_ms, "launches", st.n_launches, "fallback", st.n_fallback)

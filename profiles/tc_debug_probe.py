"""Per-role cycle accounting of the tensor scan kernel (set BFK_TC_DEBUG=1): python profiles/tc_debug_probe.py [nq ...]"""
import sys, torch
sys.path.insert(0, '/root/repo')
import docqa_b200 as bfk
n, d = 1_000_000, 384
g = torch.Generator(device="cuda").manual_seed(1234)
xb = torch.randn((n, d), generator=g, device="cuda"); xb /= xb.norm(dim=1, keepdim=True)
idx = bfk.IndexFlatL2(d); idx.add(xb)
for nq in [int(a) for a in sys.argv[1:]] or [64, 1024]:
    q = torch.randn((nq, d), device="cuda"); q /= q.norm(dim=1, keepdim=True)
    st = bfk.SearchStats()
    for i in range(3):
        idx.search(q, 10, algo="tensor", stats=st)
    print(nq, "scan_ms", st.scan_ms, "fallback", st.n_fallback, file=sys.stderr)

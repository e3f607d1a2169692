import sys, numpy as np
sys.path.insert(0, '/root/repo')
import docqa_b200 as bfk
rng = np.random.default_rng(0)
n, d, nq, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
xb = rng.standard_normal((n, d)).astype(np.float32); xb /= np.linalg.norm(xb, axis=1, keepdims=True)
xq = rng.standard_normal((nq, d)).astype(np.float32); xq /= np.linalg.norm(xq, axis=1, keepdims=True)
idx = bfk.IndexFlatL2(d); idx.add(xb)
st = bfk.SearchStats()
D, I = idx.search(xq, k, algo="tensor", stats=st)
De, Ie = idx.search(xq, k, algo="exact")
print("ok", (I == Ie).mean(), st.n_fallback)

import sys, time, numpy as np
sys.path.insert(0, "/root/repo")
import docqa_b200 as bfk
idx = bfk.read_index("/root/repo/tests/golden/docqa_vector_store.faiss")
xb = idx.reconstruct_n(0, idx.ntotal)
q = xb[:500] + 0.01
for algo in ("tensor", "exact", None):
    for i in range(50): idx.search(q[i:i+1], 5, algo=algo)
    t0 = time.perf_counter()
    for i in range(500): idx.search(q[i:i+1], 5, algo=algo)
    print(algo, (time.perf_counter() - t0) / 500 * 1e6, "us")
import torch
qd = torch.from_numpy(q).cuda()
D = torch.empty((1,5), device="cuda"); I = torch.empty((1,5), dtype=torch.int64, device="cuda")
for algo in ("tensor", "exact"):
    for i in range(50): idx.search(qd[i:i+1], 5, D=D, I=I, algo=algo)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(500): idx.search(qd[i:i+1], 5, D=D, I=I, algo=algo)
    torch.cuda.synchronize()
    print("device", algo, (time.perf_counter() - t0) / 500 * 1e6, "us")

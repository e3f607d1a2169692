"""Where a single-query search over the reference's own 649-row index spends its time (BASELINE config 1).

    python profiles/c1_latency_probe.py        (BFK_TC_SPIN=0 to compare with the blocking stream wait)
"""
import ctypes
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import docqa_b200 as bfk  # noqa: E402
from docqa_b200 import _lib  # noqa: E402

idx = bfk.read_index(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                                  "docqa_vector_store.faiss"))
xb = idx.reconstruct_n(0, idx.ntotal)
q = np.ascontiguousarray(xb[:500] + 0.01, dtype=np.float32)
tag = " ".join(f"{k[4:]}={v}" for k, v in sorted(os.environ.items()) if k.startswith("BFK_")) or "defaults"
N = 2000


def loop(fn):
    for i in range(100):
        fn(i % 500)
    t0 = time.perf_counter()
    for i in range(N):
        fn(i % 500)
    return (time.perf_counter() - t0) / N * 1e6


for algo in (None, "exact", "tensor"):
    print(f"[{tag}] python API, numpy in/out, algo={algo}: {loop(lambda i: idx.search(q[i:i + 1], 5, algo=algo)):.1f} us")
# the same without the Python wrapper's allocations: preallocated outputs, direct ctypes call
L = _lib.load()
D = np.empty((1, 5), np.float32)
I = np.empty((1, 5), np.int64)
qp = [int(q[i:i + 1].ctypes.data) for i in range(500)]
Dp, Ip = int(D.ctypes.data), int(I.ctypes.data)
print(f"[{tag}] bare ctypes call (bfk_search_ex, host pointers, auto): "
      f"{loop(lambda i: L.bfk_search_ex(idx._h, qp[i], 1, 5, Dp, Ip, 0, None, 0, 0, None)):.1f} us")
# device-resident call: what the GPU and the launches alone cost
qd = torch.from_numpy(q).cuda()
Dd = torch.empty((1, 5), device="cuda")
Id = torch.empty((1, 5), dtype=torch.int64, device="cuda")
for algo in ("exact", "tensor"):
    us = loop(lambda i: idx.search(qd[i:i + 1], 5, D=Dd, I=Id, algo=algo))
    torch.cuda.synchronize()
    print(f"[{tag}] device tensors in/out (enqueue only for exact), algo={algo}: {us:.1f} us")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for i in range(200):
    idx.search(qd[i:i + 1], 5, D=Dd, I=Id, algo="exact")
e1.record()
torch.cuda.synchronize()
print(f"[{tag}] GPU time per exact search (scan + select, back to back): {e0.elapsed_time(e1) / 200 * 1e3:.1f} us")

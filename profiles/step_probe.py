"""Step / scan timing of the tensor path on config 2 (1M x 384 fp32, L2, k=10), for A/B runs under
BFK_TC_PDL / BFK_TC_PREPASS: python profiles/step_probe.py [nq ...]

Per batch size: device step time (CUDA events around 50 searches), scan time from the in-kernel device-clock
stamps of the same 50 searches, scan time from CUDA events around the launch (separate calls, serial launches),
and the host-buffer (numpy) step time.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import docqa_b200 as bfk  # noqa: E402

n, d, k = int(os.environ.get("PROBE_ROWS", 1_000_000)), 384, 10
g = torch.Generator(device="cuda").manual_seed(1234)
xb = torch.randn((n, d), generator=g, device="cuda")
xb /= xb.norm(dim=1, keepdim=True)
idx = bfk.IndexFlatL2(d)
idx.add(xb)
tag = " ".join(f"{k[7:]}={v}" for k, v in sorted(os.environ.items()) if k.startswith("BFK_TC_")) or "defaults"
for nq in [int(a) for a in sys.argv[1:]] or [1, 64, 1024, 4096]:
    q = torch.randn((nq, d), generator=g, device="cuda")
    q /= q.norm(dim=1, keepdim=True)
    D = torch.empty((nq, k), device="cuda")
    I = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    reps = 50 if nq <= 1024 else 10
    for _ in range(5):
        idx.search(q, k, D=D, I=I, algo="tensor")
    torch.cuda.synchronize()
    idx.profile_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        idx.search(q, k, D=D, I=I, algo="tensor")
    e1.record()
    torch.cuda.synchronize()
    step = e0.elapsed_time(e1) / reps
    prof = idx.profile_end(detail=True)
    ts = idx.profile_stamps().astype(np.int64)
    lead = float(np.mean(ts[:, 0] - ts[:, 2])) * 1e-3   # prep start -> scan streaming
    trail = float(np.mean(ts[:, 3] - ts[:, 1])) * 1e-3  # scan end -> last K5 CTA out
    scan_sum, n_scans = prof["scan_ms"], prof["n"]
    st = bfk.SearchStats()
    ev = []
    for _ in range(10):
        idx.search(q, k, D=D, I=I, algo="tensor", stats=st)
        ev.append(st.scan_ms)
    qn = q.cpu().numpy()
    Dn, In = np.empty((nq, k), np.float32), np.empty((nq, k), np.int64)
    for _ in range(5):
        idx.search(qn, k, D=Dn, I=In, algo="tensor")
    e0.record()
    for _ in range(reps):
        idx.search(qn, k, D=Dn, I=In, algo="tensor")
    e1.record()
    torch.cuda.synchronize()
    host = e0.elapsed_time(e1) / reps
    # the same from page-locked buffers (what bench.py's e2e uses): the kernels address them directly
    qp = torch.empty((nq, d), dtype=torch.float32, pin_memory=True)
    qp.copy_(q)
    Dp = torch.empty((nq, k), dtype=torch.float32, pin_memory=True).numpy()
    Ip = torch.empty((nq, k), dtype=torch.int64, pin_memory=True).numpy()
    qpn = qp.numpy()
    for _ in range(5):
        idx.search(qpn, k, D=Dp, I=Ip, algo="tensor")
    e0.record()
    for _ in range(reps):
        idx.search(qpn, k, D=Dp, I=Ip, algo="tensor")
    e1.record()
    torch.cuda.synchronize()
    host_pin = e0.elapsed_time(e1) / reps
    assert (Ip == In).all() and (Ip == I.cpu().numpy()).all()
    print(f"[{tag}] nq={nq}: step {step * 1e3:.1f} us = pass {prof['pass_ms'] / reps * 1e3:.1f} (prep..rerank on the device clock) + "
          f"idle {prof['idle_ms'] / max(reps - 1, 1) * 1e3:.1f} (rerank end -> next prep start); prep->scan {lead:.1f}, scan end->K5 end {trail:.1f}; scan(stamps) {scan_sum / reps * 1e3:.1f} us over {n_scans / reps:.1f} launches, "
          f"scan(events, serial) {np.mean(ev) * 1e3:.1f} us, host step {host * 1e3:.1f} us (pageable) / {host_pin * 1e3:.1f} us (pinned), fallback {st.n_fallback}", flush=True)

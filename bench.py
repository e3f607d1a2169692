#!/usr/bin/env python
"""Benchmark of the flat-kNN hot path (BASELINE.json metric: kNN queries/sec at k=10).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--nq Q] [--workload c1|c1i|c2|c3|c4|c4e|c5]
                    [--impl ours|reference]

The head line is the SAME workload at every N (N = 1 included): BASELINE config 5 -- a fixed
100M x 384 bf16-stored corpus, IndexFlatL2, k = 10, a batch of `nq` = 64 queries per step, the rows
split row-wise over the N ranks (strong scaling; at N = 1 the whole 77 GB corpus sits on one GPU).
At N = 1 the line also carries, as extra keys, the other BASELINE configs measured in the same run:
`c2` (1M x 384 fp32, batches 1/64/1024/4096: the configuration the metric is quoted on), `c3`
(10M x 768 bf16-stored, IndexFlatIP, k = 100, batch 4096), `c4` (pool + normalise + cast + add),
`c4e` (encoder -> pool -> add), `c1` (the reference's own 649-row index, single-query latency) and `c1i` (the
reference's ingestion pattern: one add per chunk, one save per message), each with its own `roofline` / `e2e` /
`cpu_baseline`.

Prints ONE JSON line (rank 0).  `value` = queries/s with queries and results resident in HBM;
`e2e` = the same through the public numpy API with host buffers (H2D + D2H inside the timed
region); `roofline` = algorithmic bytes (or flop) of the scan kernel / its launch time, where the
launch time comes from the very steps that produced `value` (the kernel stamps the device clock
itself, see bfk_profile_begin in include/bfk.h), against MEASURED_PEAKS.json; `roofline.traffic`
is read from profiles/ncu_traffic.json (DRAM bytes of one `ncu --set full` capture of that launch);
`parity` compares the step's answers with an independent route (fp32-accumulating exact scan of each
shard -> NCCL all-gather -> merge kernel) and `result_sha` hashes the ids of the seeded batch -- the
corpus is the same stream at every N, so the hash must not depend on N.
`--impl reference` times the CPU restatement of the reference path (faiss-cpu itself is not
installable here) on the same config.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

if "--impl" in sys.argv and "reference" in sys.argv:
    # torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm is meant to use the whole box
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(os.cpu_count() or 1)
os.environ.setdefault("OMP_WAIT_POLICY", "passive")
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

CORPUS_SEED, QUERY_SEED = 1234, 4321
SPEC_HBM_GBS = 8000.0      # B200 data sheet; the north star quotes "~8 TB/s"
SPEC_BF16_TFLOPS = 2250.0  # dense bf16
L2_POLICY = "every step streams the whole corpus shard (>= 0.77 GB per GPU), far above the 126 MB L2; no flush needed"

# BASELINE.json configs that are searches: name -> (rows, d, metric (1 = L2, 0 = IP), k, storage, default batch)
SEARCH_WORKLOADS = {
    "c2": (1_000_000, 384, 1, 10, "f32", 64),
    "c3": (10_000_000, 768, 0, 100, "bf16", 4096),
    "c5": (100_000_000, 384, 1, 10, "bf16", 64),
}


def workload_string(name, n_total, d, metric, k, storage, nq):
    """The one description of a search workload both arms print (the driver compares the `config` dicts)."""
    return (f"{name}: {n_total}x{d} {'fp32' if storage == 'f32' else 'bf16-stored'} corpus, "
            f"{'IndexFlatL2' if metric == 1 else 'IndexFlatIP'}, k={k}, nq={nq}")


def search_config(name, n_total, d, metric, k, storage, nq):
    return {"workload": workload_string(name, n_total, d, metric, k, storage, nq), "l2_policy": L2_POLICY}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return {"hbm": float(j["hbm_gbs"]), "tf": float(j.get("bf16_tflops", 1590.0)),
                "tf_sustained": float(j.get("bf16_tflops_sustained", 0.0)) or None, "kind": "measured"}
    return {"hbm": 6650.0, "tf": 1590.0, "tf_sustained": None, "kind": "fallback"}


def ncu_traffic(key):
    """DRAM bytes per launch of one `ncu --set full` capture, committed under profiles/ (None when not captured)."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        ent = json.load(open(p)).get(key)
    except (OSError, ValueError):
        return None, None
    if not ent:
        return None, None
    return float(ent["dram_bytes"]), ent.get("source")


class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.FIELDS}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[3:7]):
                if v.strip().lower() == "active":
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def gen_unit_rows(torch, n, d, seed, device, chunk=1 << 20):
    """Seeded N(0, I) rows normalised to unit length, generated on the device in chunks."""
    g = torch.Generator(device=device).manual_seed(seed)
    for r0 in range(0, n, chunk):
        m = min(chunk, n - r0)
        x = torch.randn((m, d), generator=g, device=device, dtype=torch.float32)
        x /= x.norm(dim=1, keepdim=True)
        yield r0, x


def scan_roofline(res, nq, n_rows, d, k, metric, elem, pk):
    """Roofline of the dominant (scan) kernel for one measured batch size.

    Algorithmic work per launch follows SURVEY 8(d) / DESIGN.md section 4:
      exact SIMT scan (<= 8 queries per corpus pass): N*d*s bytes + queries + per-CTA lists
      tensor scan, HBM regime (nq <= 128): N*ld16*2 (bf16 copy) + N*4 (norms, L2) + nq*d*2 + lists
      tensor scan, tensor regime (nq > 128): 2*nq*N*d flop per step (all passes of the step together)
    `scan_ms` is the scan time per step and `scans_per_step` the number of scan launches in it, both from the
    timed steps themselves.
    """
    scans = max(1, int(round(res.get("scans_per_step", 1))))
    if res["algo"] == "exact_simt":
        per_q = min(nq, 8)
        bytes_launch = n_rows * d * elem + per_q * d * 4 + per_q * k * 8 * 148
        launch_ms = res["scan_ms"] / scans
        ach = bytes_launch / (launch_ms * 1e-3) / 1e9 if launch_ms > 0 else 0.0
        traffic, src = ncu_traffic(f"scan_exact_kernel|n={n_rows}|d={d}|nq={nq}|elem={elem}")
        r = {"bound": "hbm", "kernel": "scan_exact_kernel", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s",
             "frac": ach / pk["hbm"], "frac_of_spec": ach / SPEC_HBM_GBS, "peak_kind": pk["kind"] + " copy bandwidth",
             "bytes_per_launch": bytes_launch, "launch_ms": launch_ms, "launches_per_step": scans,
             "traffic": traffic, "traffic_source": src}
        return r
    ld16 = -(-d // 64) * 64
    if nq <= 128:
        launch_ms = res["scan_ms"] / scans
        bytes_launch = n_rows * ld16 * 2 + (n_rows * 4 if metric == 1 else 0) + nq * d * 2 + nq * 148 * 16 * 12
        ach = bytes_launch / (launch_ms * 1e-3) / 1e9 if launch_ms > 0 else 0.0
        traffic, src = ncu_traffic(f"scan_tc_kernel|n={n_rows}|d={d}|nq={nq}|metric={metric}")
        r = {"bound": "hbm", "kernel": "scan_tc_kernel", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s",
             "frac": ach / pk["hbm"], "frac_of_spec": ach / SPEC_HBM_GBS, "peak_kind": pk["kind"] + " copy bandwidth",
             "bytes_per_launch": bytes_launch, "launch_ms": launch_ms, "launches_per_step": scans,
             "traffic": traffic, "traffic_source": src}
        if r["frac"] > 1.0:
            # the measured peak is a COPY (half reads, half writes); a read-only stream can run above it
            r["note"] = "read-only scan above the measured copy (read+write) bandwidth"
        return r
    flops = 2.0 * nq * n_rows * d
    ach = flops / (res["scan_ms"] * 1e-3) / 1e12 if res["scan_ms"] > 0 else 0.0
    traffic, src = ncu_traffic(f"scan_tc_kernel|n={n_rows}|d={d}|nq={nq}|metric={metric}")
    r = {"bound": "tensor", "kernel": "scan_tc_kernel", "achieved": ach, "peak": pk["tf"], "unit": "TFLOP/s",
         "frac": ach / pk["tf"], "frac_of_spec": ach / SPEC_BF16_TFLOPS, "peak_kind": pk["kind"] + " burst bf16 GEMM",
         "flops_per_step": flops, "scan_ms_per_step": res["scan_ms"], "launches_per_step": scans,
         "traffic": traffic, "traffic_source": src}
    if pk["tf_sustained"]:
        r["frac_of_sustained"] = ach / pk["tf_sustained"]
    return r


# ---------------------------------------------------------------------------------------------------
# CPU legs (the oracle; the only places bench.py touches oracle/)
# ---------------------------------------------------------------------------------------------------
def cpu_knn_baseline(xb_h, xq_h, k, metric, n_full, budget_s=10.0, max_reps=50):
    """The faiss restatement on this box's cores over a bounded sample; value scaled linearly to n_full rows."""
    from oracle import cpu_baseline

    cpu_baseline.build()
    nq = xq_h.shape[0]
    bn = np.einsum("ij,ij->i", xb_h, xb_h, dtype=np.float32)
    fn = (lambda: cpu_baseline.knn_seq(xq_h, xb_h, k, metric)) if nq < 20 else (
        lambda: cpu_baseline.knn_blas(xq_h, xb_h, k, metric, bn=bn))
    fn()
    t0 = time.perf_counter()
    reps = 0
    while time.perf_counter() - t0 < budget_s and reps < max_reps:
        fn()
        reps += 1
    dt = time.perf_counter() - t0
    cores = min(nq, cpu_baseline.num_threads()) if nq < 20 else cpu_baseline.num_threads()
    n_s = xb_h.shape[0]
    sample = (f"{reps} calls of nq={nq} over {n_s}x{xb_h.shape[1]} fp32 rows ({'seq' if nq < 20 else 'blas'} path of "
              f"the faiss restatement)")
    if n_s != n_full:
        sample += f"; a {n_s}-row slice of the {n_full}-row corpus, value scaled linearly in N"
    return {"value": nq * reps / dt * (n_s / n_full), "unit": "queries/s", "cores": cores, "kind": "port",
            "host_cpus": os.cpu_count(), "sample": sample, "extrapolated": n_s != n_full}


def run_reference(args):
    """CPU arm: the oracle's restatement of faiss-cpu IndexFlat.search on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cpu_baseline

    cpu_baseline.build()
    name = args.workload or "c5"
    if name not in SEARCH_WORKLOADS:
        print(json.dumps({"impl": "reference", "unavailable": f"the CPU arm times the search workloads (c2, c3, c5), not {name}"}))
        return
    n_full, d, metric, k, storage, nq_def = SEARCH_WORKLOADS[name]
    nq = args.nq or nq_def
    if args.rows:
        n_full = args.rows
    # a step must stay a bounded sample: at most 1M rows x 384 (or the same bytes at another width) per step
    n_sample = min(n_full, max(1, 384_000_000 // d) if nq <= 64 else max(1, 76_800_000 // d))
    rng = np.random.default_rng(CORPUS_SEED)
    xb = rng.standard_normal((n_sample, d), dtype=np.float32)
    xb /= np.linalg.norm(xb, axis=1, keepdims=True)
    xq = np.random.default_rng(QUERY_SEED).standard_normal((nq, d), dtype=np.float32)
    xq /= np.linalg.norm(xq, axis=1, keepdims=True)
    bn = np.einsum("ij,ij->i", xb, xb, dtype=np.float32)

    # a real faiss-cpu is preferred the day one is importable (none is in this image)
    real = None
    try:
        import faiss as _faiss

        if getattr(_faiss, "__version__", "bfk-1.0") != "bfk-1.0" and hasattr(_faiss, "omp_get_max_threads"):
            real = _faiss.IndexFlatL2(d) if metric == 1 else _faiss.IndexFlatIP(d)
            real.add(xb)
    except Exception:
        real = None

    def step():
        if real is not None:
            real.search(xq, k)
        elif nq < 20:
            cpu_baseline.knn_seq(xq, xb, k, metric)
        else:
            cpu_baseline.knn_blas(xq, xb, k, metric, bn=bn)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    scale = n_sample / n_full  # brute force is linear in N
    qps = nq * args.steps / dt * scale
    cores = min(nq, cpu_baseline.num_threads()) if nq < 20 else cpu_baseline.num_threads()
    kind, ref_kind = "port", "CPU restatement of faiss-cpu IndexFlat.search (faiss not installable here)"
    if real is not None:
        cores, kind, ref_kind = _faiss.omp_get_max_threads(), "reference", f"faiss-cpu {_faiss.__version__}"
    sample = (f"full workload: {n_sample}x{d} fp32 corpus, nq={nq}" if scale == 1.0 else
              f"each step searches a {n_sample}-row slice of the {n_full}-row corpus with nq={nq} (ms_per_step is that "
              f"measured slice time); value = slice queries/s x {scale:g}, brute force being linear in N")
    line = {
        "metric": f"kNN queries/sec at k={k}", "value": qps, "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "impl": "reference", "reference_kind": ref_kind, "extrapolated": scale != 1.0,
        "config": search_config(name, n_full, d, metric, k, storage, nq),
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": kind, "sample": sample,
                         "host_cpus": os.cpu_count()},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------
# side workloads (N = 1): each returns a dict shaped like a bench line
# ---------------------------------------------------------------------------------------------------
def run_c4(args, with_cpu=True):
    """BASELINE config 4: encoder outputs [B, 256, 384] -> fused pool + normalise + cast -> index.add.

    One step = one add_pooled launch over a device batch of 16384 chunks, 62 steps = 1 015 808 chunks appended to
    one live index; 4 rotating input buffers (3.2 GB each in bf16, so every launch streams from HBM, far above the
    126 MB L2).
    """
    import torch

    import docqa_b200 as bfk

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    B, T, H = args.rows or 16384, 256, 384
    pk = peaks()
    out = {}
    steps = max(args.steps, 62)  # 62 launches of 16 384 chunks ~ the 1M chunks of BASELINE config 4
    e2e = None
    for name, dt, sz in (("bf16", torch.bfloat16, 2), ("f32", torch.float32, 4)):
        g = torch.Generator(device=dev).manual_seed(CORPUS_SEED)
        bufs = [torch.randn((B, T, H), generator=g, device=dev, dtype=torch.float32).to(dt) for _ in range(4)]
        for label, lengths in (("full", None), ("ragged", "U16..256")):
            if lengths is None:
                mask = torch.ones((B, T), dtype=torch.int64, device=dev)
            else:
                ln = torch.randint(16, T + 1, (B,), generator=g, device=dev)
                mask = (torch.arange(T, device=dev)[None, :] < ln[:, None]).to(torch.int64)
            index = bfk.IndexFlatL2(H, device=0)
            index.reserve(B * (steps + args.warmup) + 16)
            for i in range(args.warmup):
                index.add_pooled(bufs[i % 4], mask)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                index.add_pooled(bufs[i % 4], mask)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            assert index.ntotal == B * (steps + args.warmup)
            valid = float(mask.sum().item()) / B
            bytes_chunk = valid * H * sz + T * 8 + H * 4 + H * 2 + 4
            ach = bytes_chunk * B / (ms * 1e-3) / 1e9
            out[f"{name}_{label}"] = {"chunks_per_s": B / (ms * 1e-3), "ms_per_launch": ms, "valid_tokens": valid,
                                      "bytes_per_chunk": bytes_chunk, "achieved_gbs": ach, "frac": ach / pk["hbm"],
                                      "frac_of_spec": ach / SPEC_HBM_GBS}
            if name == "bf16" and label == "full":
                # end to end: token states and mask start in pinned host memory (a 1024-chunk slice per step: the
                # full 3.2 GB batch would only time PCIe), the ntotal read-back closes the step
                Be = min(B, 1024)
                tok_pin = bufs[0][:Be].cpu().pin_memory()
                mask_pin = mask[:Be].cpu().pin_memory()
                idx2 = bfk.IndexFlatL2(H, device=0)
                e_steps = min(steps, 10)
                idx2.reserve(Be * (e_steps + 2))
                t0 = time.perf_counter()
                for i in range(e_steps):
                    idx2.add_pooled(tok_pin.to(dev, non_blocking=True), mask_pin.to(dev, non_blocking=True))
                    torch.cuda.synchronize()
                    assert idx2.ntotal == Be * (i + 1)
                ms_e = (time.perf_counter() - t0) * 1e3 / e_steps
                e2e = {"value": Be / (ms_e * 1e-3), "unit": "chunks/s", "h2d_bytes_per_step": Be * T * (H * sz + 8),
                       "d2h_bytes_per_step": 8, "chunks_per_step": Be,
                       "note": "PCIe-bound: 197 KB of token states per chunk cross the bus"}
                del idx2
            del index
        del bufs
        torch.cuda.empty_cache()
    # K2: index.add of fp32 rows already on the device (indexer.py:41 at scale): 4d bytes in, fp32 master row +
    # bf16 scan row + norm out
    rows_n = 4_000_000
    g = torch.Generator(device=dev).manual_seed(CORPUS_SEED)
    xb = torch.randn((rows_n, H), generator=g, device=dev, dtype=torch.float32)
    for storage in ("f32", "bf16"):
        index = bfk.IndexFlatL2(H, device=0, storage=storage)
        reps = 6
        index.reserve(rows_n * (reps + 2))
        index.add(xb)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            index.add(xb)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        bytes_row = H * 4 + (H * 4 if storage == "f32" else 0) + H * 2 + 4
        ach = bytes_row * rows_n / (ms * 1e-3) / 1e9
        out[f"add_rows_{storage}"] = {"rows_per_s": rows_n / (ms * 1e-3), "ms_per_launch": ms, "bytes_per_row": bytes_row,
                                      "achieved_gbs": ach, "frac": ach / pk["hbm"], "kernel": "append_rows_kernel"}
        del index
    del xb
    torch.cuda.empty_cache()
    cpu = None
    if with_cpu:
        from oracle import pool_ref

        rng = np.random.default_rng(CORPUS_SEED)
        Bc = 256
        tok = rng.standard_normal((Bc, T, H), dtype=np.float32)
        msk = np.ones((Bc, T), dtype=np.int64)
        pool_ref.mean_pool_normalize(tok[:8], msk[:8])
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 5.0 and reps < 20:
            pool_ref.mean_pool_normalize(tok, msk)
            reps += 1
        dt = time.perf_counter() - t0
        cpu = {"value": Bc * reps / dt, "unit": "chunks/s", "cores": 1, "kind": "port",
               "sample": f"{reps} calls of the numpy restatement of Pooling(mean)+Normalize over {Bc} chunks x {T} x {H} fp32 "
                         "(no index append)"}
    head = out["bf16_full"]
    traffic, src = ncu_traffic(f"pool_stream_kernel|B={B}|T={T}|H={H}|bf16|full")
    return {
        "metric": "ingestion chunks/sec (pool+normalise+cast+add)", "value": head["chunks_per_s"], "unit": "chunks/s",
        "n_gpus": 1, "steps": steps, "warmup": args.warmup, "ms_per_step": head["ms_per_launch"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16 in, f32 accumulate", "data": "synthetic",
        "config": {"workload": f"c4: token states [{B}, {T}, {H}] bf16, int64 mask, fused pool+normalise+cast+index.add",
                   "l2_policy": "4 rotating 3.2 GB input buffers, far above the 126 MB L2"},
        "roofline": {"bound": "hbm", "kernel": "pool_stream_kernel", "achieved": head["achieved_gbs"], "peak": pk["hbm"],
                     "unit": "GB/s", "frac": head["frac"], "frac_of_spec": head["frac_of_spec"],
                     "peak_kind": pk["kind"] + " copy bandwidth", "bytes_per_launch": head["bytes_per_chunk"] * B,
                     "launch_ms": head["ms_per_launch"], "traffic": traffic, "traffic_source": src},
        "e2e": e2e, "cpu_baseline": cpu, "gpu_launches": steps, "variants": out,
    }


def run_c1(args):
    """BASELINE config 1: the reference's own 649 x 384 index, one query per call, k = 5 (llm-qa/main.py:101,117).

    Too small for a roofline ('parity + latency only'): the number is the wall-clock latency of one
    `index.search(np.float32[1, 384], 5)` through the numpy API, host buffers in and out, next to the CPU
    restatement of faiss on the same box.
    """
    import docqa_b200 as bfk
    from oracle import cpu_baseline, faiss_io

    path = os.path.join(ROOT, "tests", "golden", "docqa_vector_store.faiss")
    index = bfk.read_index(path)
    _, n, metric, xb = faiss_io.read_index_bytes(open(path, "rb").read())
    rng = np.random.default_rng(QUERY_SEED)
    q = xb[rng.integers(0, n, 2000)] + 0.05 * rng.standard_normal((2000, xb.shape[1])).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    for i in range(50):
        index.search(q[i : i + 1], 5)
    lat = []
    for i in range(2000):
        t0 = time.perf_counter()
        index.search(q[i : i + 1], 5)
        lat.append(time.perf_counter() - t0)
    lat = np.sort(np.array(lat)) * 1e3
    cpu_baseline.build()
    cl = []
    for i in range(300):
        t0 = time.perf_counter()
        cpu_baseline.knn_seq(q[i : i + 1], xb, 5, metric)
        cl.append(time.perf_counter() - t0)
    cl = np.sort(np.array(cl)) * 1e3
    return {
        "metric": "single-query search latency at k=5 (semantic-indexer default_data)", "value": float(np.median(lat)),
        "unit": "ms", "n_gpus": 1, "steps": 2000, "warmup": 50, "ms_per_step": float(np.median(lat)),
        "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "the reference's shipped vector_store.faiss (649 x 384)",
        "config": {"workload": "c1: 649x384 fixture, IndexFlatL2, k=5, nq=1 per call, numpy in / numpy out"},
        "latency_ms": {"p50": float(np.median(lat)), "p90": float(lat[int(0.9 * len(lat))]), "p99": float(lat[int(0.99 * len(lat))])},
        "e2e": {"value": float(np.median(lat)), "unit": "ms", "h2d_bytes_per_step": 384 * 4, "d2h_bytes_per_step": 5 * 12},
        "gpu_launches": 2 * 2000, "roofline": None,
        "cpu_baseline": {"value": float(np.median(cl)), "unit": "ms", "cores": 1, "kind": "port",
                         "sample": "300 single-query calls of the faiss restatement (seq path) on the same 649 rows"},
    }


def run_c1i(args):
    """The reference's own ingestion pattern, timed (indexer.py:37-41,120-125): per message, every chunk is added
    on its own (`index.add(np.float32[1, 384])`) and the state is saved once.

    1000 messages x 3 chunks into an index that starts from the reference's 649-row file.  Three variants on the
    GPU index: the reference's call shape with a full rewrite per message (save_state as indexer.py:26-30 does
    it), the same with the append-only save, and one batched add per message with the append-only save; next to
    the CPU restatement (numpy append + full rewrite of the faiss file and the pickle).
    """
    import pickle
    import shutil

    import docqa_b200 as bfk
    from docqa_b200 import indexer_host
    from oracle import faiss_io

    src = os.path.join(ROOT, "tests", "golden", "docqa_vector_store.faiss")
    _, n0, metric, xb0 = faiss_io.read_index_bytes(open(src, "rb").read())
    n_msg, per_msg, d = args.rows or 1000, 3, xb0.shape[1]
    rng = np.random.default_rng(CORPUS_SEED)
    vec = rng.standard_normal((n_msg * per_msg, d)).astype(np.float32)
    vec /= np.linalg.norm(vec, axis=1, keepdims=True)
    out = {}
    tmp = tempfile.mkdtemp(prefix="bfk_c1i_")
    try:
        for name, batched, incremental in (("n1_adds_full_rewrite", False, False), ("n1_adds_append_only", False, True),
                                           ("batched_add_append_only", True, True)):
            wd = os.path.join(tmp, name)
            os.makedirs(wd)
            shutil.copy(src, os.path.join(wd, "vector_store.faiss"))
            meta0 = [{"doc_id": "KB_MTC", "text_content": f"row {i}", "source": "fixture", "type": "knowledge_base"} for i in range(n0)]
            with open(os.path.join(wd, "metadata_store.pkl"), "wb") as f:
                pickle.dump(meta0, f)
            ix = indexer_host.SemanticIndexer(None, d, index_file=os.path.join(wd, "vector_store.faiss"),
                                              metadata_file=os.path.join(wd, "metadata_store.pkl"),
                                              incremental_save=incremental)
            ix.load_or_build()
            lat = []
            for m in range(n_msg):
                rows = vec[m * per_msg : (m + 1) * per_msg]
                metas = [{"doc_id": "KB_MTC", "text_content": f"msg {m} chunk {j} " + "x" * 480, "source": f"Dossier Patient {m}",
                          "type": "patient_file"} for j in range(per_msg)]
                t0 = time.perf_counter()
                if batched:
                    ix.add_vectors(rows, metas)
                else:
                    for j in range(per_msg):
                        ix.add_vectors(rows[j : j + 1], metas[j : j + 1])
                ix.save_state()
                lat.append(time.perf_counter() - t0)
            lat = np.sort(np.array(lat)) * 1e3
            assert ix.index.ntotal == n0 + n_msg * per_msg
            out[name] = {"ms_per_message_p50": float(np.median(lat)), "p90": float(lat[int(0.9 * len(lat))]),
                         "messages_per_s": float(n_msg / (lat.sum() * 1e-3))}
        # CPU restatement of indexer.py: numpy rows, full rewrite of both files per message
        xb = xb0.copy()
        meta = [{"doc_id": "KB_MTC", "text_content": f"row {i}", "source": "fixture", "type": "knowledge_base"} for i in range(n0)]
        wd = os.path.join(tmp, "cpu")
        os.makedirs(wd)
        lat = []
        for m in range(min(n_msg, 300)):
            rows = vec[m * per_msg : (m + 1) * per_msg]
            t0 = time.perf_counter()
            for j in range(per_msg):
                xb = np.concatenate([xb, rows[j : j + 1]], axis=0)
                meta.append({"doc_id": "KB_MTC", "text_content": f"msg {m} chunk {j} " + "x" * 480, "source": f"Dossier Patient {m}",
                             "type": "patient_file"})
            with open(os.path.join(wd, "vector_store.faiss"), "wb") as f:
                f.write(faiss_io.write_index_bytes(xb, metric))
            with open(os.path.join(wd, "metadata_store.pkl"), "wb") as f:
                pickle.dump(meta, f)
            lat.append(time.perf_counter() - t0)
        lat = np.sort(np.array(lat)) * 1e3
        cpu = {"value": float(np.median(lat)), "unit": "ms", "cores": 1, "kind": "port",
               "sample": f"{len(lat)} messages x {per_msg} chunks: numpy append + full rewrite of the faiss file and the pickle"}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    head = out["n1_adds_append_only"]
    return {
        "metric": "ingestion latency per message (3 chunks: n=1 adds + save_state)", "value": head["ms_per_message_p50"],
        "unit": "ms", "n_gpus": 1, "steps": n_msg, "warmup": 0, "ms_per_step": head["ms_per_message_p50"],
        "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic unit rows appended to the reference's 649-row index",
        "config": {"workload": f"c1i: {n_msg} messages x {per_msg} chunks, index.add(1x384) per chunk + save_state per message"},
        "e2e": {"value": head["ms_per_message_p50"], "unit": "ms", "h2d_bytes_per_step": per_msg * d * 4,
                "d2h_bytes_per_step": per_msg * d * 4},
        "gpu_launches": n_msg * per_msg, "roofline": None, "cpu_baseline": cpu, "variants": out,
    }


def run_c4e(args):
    """Config 4 as a real pipeline: token ids -> 6-layer encoder -> pool + normalise -> index.add.

    One step = encode_add of B chunks of 256 tokens (random ids, random-initialised all-MiniLM-L6-v2-shaped
    weights: the checkpoint is not available offline).  Work per chunk: 6 layers x 2*T*(3H*H + H*H + 2*H*I)
    GEMM flop + 6 x 4*T*T*H attention flop = 5.6 + 0.6 GFLOP at T=256.
    """
    import torch

    import docqa_b200 as bfk

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    B, T, H, I, L = args.rows or 2048, 256, 384, 1536, 6
    pk = peaks()
    enc = bfk.MiniLMEncoder(device=0)
    g = torch.Generator().manual_seed(CORPUS_SEED)
    cfg = enc.config
    sd = {"embeddings.word_embeddings.weight": 0.05 * torch.randn((cfg["vocab_size"], H), generator=g),
          "embeddings.position_embeddings.weight": 0.05 * torch.randn((cfg["max_position_embeddings"], H), generator=g),
          "embeddings.token_type_embeddings.weight": 0.05 * torch.randn((2, H), generator=g),
          "embeddings.LayerNorm.weight": torch.ones(H), "embeddings.LayerNorm.bias": torch.zeros(H)}
    for li in range(L):
        p = f"encoder.layer.{li}."
        for name, shape in (("attention.self.query", (H, H)), ("attention.self.key", (H, H)), ("attention.self.value", (H, H)),
                            ("attention.output.dense", (H, H)), ("intermediate.dense", (I, H)), ("output.dense", (H, I))):
            sd[p + name + ".weight"] = 0.05 * torch.randn(shape, generator=g)
            sd[p + name + ".bias"] = 0.05 * torch.randn(shape[0], generator=g)
        for name in ("attention.output.LayerNorm", "output.LayerNorm"):
            sd[p + name + ".weight"] = torch.ones(H)
            sd[p + name + ".bias"] = torch.zeros(H)
    enc.load_state_dict(sd)
    gd = torch.Generator(device=dev).manual_seed(QUERY_SEED)
    bufs = [torch.randint(0, cfg["vocab_size"], (B, T), generator=gd, device=dev) for _ in range(4)]
    mask = torch.ones((B, T), dtype=torch.int64, device=dev)
    steps = max(min(args.steps, 10), 5)
    index = bfk.IndexFlatL2(H, device=0)
    index.reserve(B * (steps + args.warmup) + 16)
    sampler = ClockSampler(0)
    for i in range(args.warmup):
        enc.encode_add(index, bufs[i % 4], mask)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        enc.encode_add(index, bufs[i % 4], mask)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    clocks = sampler.stop()
    assert index.ntotal == B * (steps + args.warmup)
    # end to end: ids and mask start in pinned host memory, the ntotal read-back closes the step
    ids_pin = bufs[0].cpu().pin_memory()
    mask_pin = mask.cpu().pin_memory()
    index2 = bfk.IndexFlatL2(H, device=0)
    index2.reserve(B * (steps + 2))
    t0 = time.perf_counter()
    for i in range(steps):
        enc.encode_add(index2, ids_pin.to(dev, non_blocking=True), mask_pin.to(dev, non_blocking=True))
        torch.cuda.synchronize()
        assert index2.ntotal == B * (i + 1)
    ms_e2e = (time.perf_counter() - t0) * 1e3 / steps
    gemm_flop = L * 2.0 * T * (3 * H * H + H * H + 2 * H * I)
    attn_flop = L * 4.0 * T * T * H
    ach = (gemm_flop + attn_flop) * B / (ms * 1e-3) / 1e12
    return {
        "metric": "ingestion chunks/sec (encoder + pool + normalise + add)", "value": B / (ms * 1e-3), "unit": "chunks/s",
        "n_gpus": 1, "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16 activations, f32 accumulate", "data": "synthetic",
        "config": {"workload": f"c4e: {B} chunks x {T} tokens, BERT 6x384 (12 heads, 1536) random weights -> mean pool -> index.add",
                   "l2_policy": "4 rotating id buffers; activations of a step (2 GB) far exceed the 126 MB L2"},
        "e2e": {"value": B / (ms_e2e * 1e-3), "unit": "chunks/s", "h2d_bytes_per_step": B * T * 16, "d2h_bytes_per_step": 8},
        # per micro-batch of 131072 tokens: embed + 5 kernels per layer + pool/append
        "gpu_launches": steps * -(-B * T // 131072) * (1 + 5 * L + 1), "clocks": clocks,
        "roofline": {"bound": "tensor", "kernel": "enc_gemm_kernel (+ enc_attention_pws_kernel)", "achieved": ach, "peak": pk["tf"],
                     "unit": "TFLOP/s", "frac": ach / pk["tf"], "frac_of_spec": ach / SPEC_BF16_TFLOPS,
                     "frac_of_sustained": (ach / pk["tf_sustained"]) if pk["tf_sustained"] else None,
                     "peak_kind": pk["kind"] + " burst bf16 GEMM", "flops_per_launch": (gemm_flop + attn_flop) * B,
                     "launch_ms": ms, "traffic": None,
                     "note": "whole forward pass (32 launches per 512-chunk micro-batch) against the tensor peak; per-kernel "
                             "split in profiles/r02_launches_encoder.csv"},
        "cpu_baseline": None,
    }


# ---------------------------------------------------------------------------------------------------
# the search workloads (c2, c3, c5), single GPU or row-sharded
# ---------------------------------------------------------------------------------------------------
class SearchBench:
    def __init__(self, args, name, torch, dist, bfk, rank, world, local_rank):
        self.args, self.name, self.torch, self.dist, self.bfk = args, name, torch, dist, bfk
        self.rank, self.world, self.local_rank = rank, world, local_rank
        self.dev = torch.device("cuda", local_rank)
        n_total, self.d, self.metric, self.k, self.storage, self.nq_def = SEARCH_WORKLOADS[name]
        self.n_total = args.rows or n_total
        self.lo, self.hi = bfk.shard_range(self.n_total, rank, world)
        self.pk = peaks()
        self.index = self.local = None

    def build(self):
        torch, bfk = self.torch, self.bfk
        if self.world > 1:
            self.index = bfk.ShardedIndexFlat(self.d, self.metric, device=self.local_rank, storage=self.storage)
            self.local = self.index.local
        else:
            self.index = self.local = bfk.IndexFlat(self.d, self.metric, device=self.local_rank, storage=self.storage)
        self.local.reserve(self.hi - self.lo)
        # every rank draws the same global stream and keeps its slice, so the corpus does not depend on N
        for r0, x in gen_unit_rows(torch, self.n_total, self.d, CORPUS_SEED, self.dev):
            a, b = max(self.lo, r0), min(self.hi, r0 + x.shape[0])
            if b > a:
                self.local.add(x[a - r0 : b - r0])
            del x
        if self.world > 1:
            self.index.set_contiguous_shard(self.lo, self.n_total)
        torch.cuda.synchronize()
        assert self.local.ntotal == self.hi - self.lo

    def free(self):
        self.index = self.local = None
        self.torch.cuda.empty_cache()

    def queries(self, nq):
        torch = self.torch
        g = torch.Generator(device=self.dev).manual_seed(QUERY_SEED)
        q = torch.randn((nq, self.d), generator=g, device=self.dev, dtype=torch.float32)
        return q / q.norm(dim=1, keepdim=True)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def time_steps(self, fn, steps, warmup, profile=False):
        """W untimed steps, then exactly K steps bracketed by barrier + synchronize; CUDA events, max over ranks.
        With profile=True the scan kernels of those K steps stamp the device clock (no event sits between the
        kernels of a search): returns (ms total, scan ms total of this rank's shard, scan launches)."""
        torch = self.torch
        for _ in range(warmup):
            fn()
        self.barrier()
        if profile:
            self.local.profile_begin()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        self.barrier()
        ms = self.max_over_ranks(e0.elapsed_time(e1))
        if not profile:
            return ms, None, None
        scan_ms, n_scans = self.local.profile_end()
        return ms, self.max_over_ranks(scan_ms), n_scans

    def measure(self, nq, steps, warmup, with_e2e=True, sustain_s=0.0):
        torch, K = self.torch, self.k
        q_dev = self.queries(nq)
        D_dev = torch.empty((nq, K), dtype=torch.float32, device=self.dev)
        I_dev = torch.empty((nq, K), dtype=torch.int64, device=self.dev)
        kw = {"algo": self.args.algo} if self.args.algo else {}
        index, local, world = self.index, self.local, self.world

        if world == 1:
            def dev_step():
                index.search(q_dev, K, D=D_dev, I=I_dev, **kw)
        else:
            def dev_step():
                index.search(q_dev, K, **kw)

        # one un-timed call with the library's own statistics: which algorithm, how many launches, fallbacks
        st = self.bfk.SearchStats()
        local.search(q_dev, K, stats=st, **kw)
        algo = {1: "exact_simt", 2: "tensor"}.get(st.algo_used, "?")
        ms, scan_ms, n_scans = self.time_steps(dev_step, steps, warmup, profile=(algo == "tensor"))
        out = {"nq": nq, "value": nq * steps / (ms * 1e-3), "ms_per_step": ms / steps, "algo": algo,
               "n_fallback": int(st.n_fallback), "launches_per_step": st.n_launches + (4 if world > 1 else 0)}
        if algo == "tensor" and n_scans:
            out["scan_ms"] = scan_ms / steps
            out["scans_per_step"] = n_scans / steps
            out["scan_ms_source"] = "in-kernel device-clock stamps of the timed steps (bfk_profile_begin/end)"
        else:
            # the fp32 scan carries no stamps: CUDA events around its launches, in separate calls
            sm = []
            for _ in range(min(steps, 10)):
                local.search(q_dev, K, stats=st, **kw)
                sm.append(st.scan_ms)
            out["scan_ms"] = statistics.mean(sm)
            out["scans_per_step"] = -(-nq // 8)
            out["scan_ms_source"] = "CUDA events around the scan launches of separate calls"
        if with_e2e:
            q_pin = torch.empty((nq, self.d), dtype=torch.float32, pin_memory=True)
            q_pin.copy_(q_dev)
            q_np = q_pin.numpy()
            if world == 1:
                D_np = torch.empty((nq, K), dtype=torch.float32, pin_memory=True).numpy()
                I_np = torch.empty((nq, K), dtype=torch.int64, pin_memory=True).numpy()

                def host_step():
                    index.search(q_np, K, D=D_np, I=I_np, **kw)
            else:
                def host_step():
                    index.search(q_np, K, **kw)

            ms_h, _, _ = self.time_steps(host_step, steps, warmup)
            out["e2e"] = {"value": nq * steps / (ms_h * 1e-3), "unit": "queries/s", "ms_per_step": ms_h / steps,
                          "h2d_bytes_per_step": nq * self.d * 4, "d2h_bytes_per_step": nq * K * 12}
        if sustain_s > 0:
            # the K timed steps last only milliseconds; keep the same step running long enough for
            # nvidia-smi (100 ms period) to see the clocks the kernel actually runs at, and report
            # the throughput of that longer window too (last, so that the device-resident and the
            # host-buffer loops above both run before the power cap has pulled the clocks down)
            n_rep = max(steps, int(sustain_s / max(ms / steps * 1e-3, 1e-6)))
            ms_s, _, _ = self.time_steps(dev_step, n_rep, 0)
            out["sustained"] = {"value": nq * n_rep / (ms_s * 1e-3), "steps": n_rep, "seconds": ms_s * 1e-3}
        out["roofline"] = scan_roofline(out, nq, self.hi - self.lo, self.d, K, self.metric,
                                        4 if self.storage == "f32" else 2, self.pk)
        return out

    def parity(self, nq_par=16, nq_sha=None):
        """The product path's answers against an independent route, compared on the GPU.

        Independent route: every rank scans its shard with the fp32-accumulating exact kernel (K3a, no tensor
        cores, no certificate, no deferred search), ids made global by `id_offset`; the per-rank lists travel
        through ONE NCCL all-gather and are merged by K6 (`bfk_merge_topk`).  It shares neither the tensor scan
        nor the peer-memory exchange with the product path.  Tolerances are the north star's: ids identical
        except inside distance ties of 1e-4 relative, distances within 1e-3 relative.
        """
        torch, K = self.torch, self.k
        nq_sha = nq_sha or self.nq_def
        q = self.queries(nq_sha)
        Dp, Ip = self.index.search(q, K)
        torch.cuda.synchronize()
        sha = hashlib.sha1(Ip.cpu().numpy().astype(np.int64).tobytes()).hexdigest()
        qs = q[:nq_par].contiguous()
        Dl, Il = self.local.search(qs, K, algo="exact", id_offset=self.lo)
        if self.world > 1:
            from docqa_b200.sharded import _cuda_merge

            Dg = torch.empty((self.world, nq_par, K), dtype=torch.float32, device=self.dev)
            Ig = torch.empty((self.world, nq_par, K), dtype=torch.int64, device=self.dev)
            self.dist.all_gather_into_tensor(Dg, Dl.contiguous())
            self.dist.all_gather_into_tensor(Ig, Il.contiguous())
            Dr, Ir = _cuda_merge(Dg, Ig, K, self.metric)
        else:
            Dr, Ir = Dl, Il
        Dq, Iq = Dp[:nq_par], Ip[:nq_par]
        scale = torch.maximum(Dr.abs(), torch.full_like(Dr, 1e-6))
        rel = ((Dq - Dr).abs() / scale)
        id_diff = Iq != Ir
        # a differing id is acceptable only inside a tie: both routes report (nearly) the same distance there
        bad_ids = int((id_diff & (rel > 1e-4)).sum().item())
        max_rel = float(rel.max().item())
        ok = bad_ids == 0 and max_rel <= 1e-3
        if self.world > 1:
            t = torch.tensor([0 if ok else 1], device=self.dev, dtype=torch.int64)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ok = int(t.item()) == 0
        return {"ok": bool(ok), "queries_compared": nq_par, "k": K, "ids_differing": int(id_diff.sum().item()),
                "ids_differing_outside_ties": bad_ids, "max_rel_distance_error": max_rel,
                "reference_route": "per-shard fp32 exact scan (K3a) -> NCCL all_gather_into_tensor -> bfk_merge_topk"
                                   if self.world > 1 else "fp32 exact scan (K3a + K4) of the same index",
                "tolerance": "ids equal except inside ties <= 1e-4 relative; distances <= 1e-3 relative"}, sha

    def cpu_baseline(self, nq, budget_s=10.0):
        # a bounded sample: 1M rows for the small batches, 100k rows for the 4096-query batches (0.6 TFLOP per call)
        rows = 1_000_000 if nq <= 64 else 100_000
        xb_h = self.local.reconstruct_n(0, min(self.hi - self.lo, rows))
        xq_h = self.queries(nq).cpu().numpy()
        return cpu_knn_baseline(xb_h, xq_h, self.k, self.metric, self.n_total, budget_s=budget_s)


def search_line(sb, args, nq, batches, with_cpu, clocks_for_head=True):
    """One bench line (dict) for a search workload: head batch `nq` + the other batch sizes under `batches`."""
    sampler = ClockSampler(sb.local_rank) if (sb.rank == 0 and clocks_for_head) else None
    if sampler:
        time.sleep(0.3)  # let nvidia-smi start sampling
    head = sb.measure(nq, args.steps, args.warmup, sustain_s=1.5 if clocks_for_head else 0.0)
    clocks = sampler.stop() if sampler else None
    extra = {}
    for b in batches:
        if b != nq:
            r = sb.measure(b, max(3, min(args.steps, 10)), 3, with_e2e=True)
            extra[str(b)] = {"value": r["value"], "ms_per_step": r["ms_per_step"], "algo": r["algo"],
                             "e2e": r["e2e"]["value"], "e2e_ms_per_step": r["e2e"]["ms_per_step"], "scan_ms": r["scan_ms"],
                             "scans_per_step": r["scans_per_step"], "n_fallback": r["n_fallback"], "roofline": r["roofline"]}
    par, sha = sb.parity(nq_par=min(16, nq), nq_sha=nq)
    cpu = sb.cpu_baseline(nq) if (with_cpu and sb.rank == 0 and sb.world == 1) else None
    cfg = search_config(sb.name, sb.n_total, sb.d, sb.metric, sb.k, sb.storage, nq)
    line = {
        "metric": f"kNN queries/sec at k={sb.k}", "value": head["value"], "unit": "queries/s", "n_gpus": sb.world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None,
        "dtype": "f32" if head["algo"] == "exact_simt" else "bf16 scan + f32 rerank", "data": "synthetic",
        "config": cfg, "algo": head["algo"], "sharding": f"rows split over {sb.world} GPUs" if sb.world > 1 else "one GPU",
        "e2e": head["e2e"], "gpu_launches": head["launches_per_step"] * args.steps, "clocks": clocks,
        "scan_ms": head["scan_ms"], "scans_per_step": head["scans_per_step"], "scan_ms_source": head["scan_ms_source"],
        "roofline": head["roofline"], "cpu_baseline": cpu, "batches": extra, "n_fallback": head["n_fallback"],
        "sustained": head.get("sustained"), "parity": par, "result_sha": sha,
    }
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--nq", type=int, default=None)
    ap.add_argument("--workload", choices=["c1", "c1i", "c2", "c3", "c5", "c4", "c4e"], default=None)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--rows", type=int, default=None, help="override corpus rows (debugging)")
    ap.add_argument("--algo", default=None, choices=[None, "auto", "exact", "tensor"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="N = 1: skip the c2 / c3 / c4 / c4e / c1 extra keys")
    ap.add_argument("--no-scaling-base", action="store_true", help="(kept for old command lines; same as --no-extras)")
    ap.add_argument("--batches", default=None, help="extra batch sizes reported under `batches`")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)
    side = {"c1": run_c1, "c1i": run_c1i, "c4": run_c4, "c4e": run_c4e}
    if args.workload in side:
        import torch

        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
        print(json.dumps(side[args.workload](args)))
        return

    import torch
    import torch.distributed as dist

    import docqa_b200 as bfk

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if world != args.gpus:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}", file=sys.stderr)

    name = args.workload or "c5"
    default_batches = {"c2": [1, 64, 1024, 4096], "c3": [], "c5": [1, 64, 1024, 8192]}[name]
    batches = [int(s) for s in args.batches.split(",") if s] if args.batches is not None else default_batches
    nq = args.nq or SEARCH_WORKLOADS[name][5]
    with_cpu = not args.no_cpu_baseline

    sb = SearchBench(args, name, torch, dist, bfk, rank, world, local_rank)
    if world == 1 and name == "c5" and args.rows is None:
        free_b, _ = torch.cuda.mem_get_info(dev)
        need = sb.n_total * (sb.d * 2 + 4) + (8 << 30)
        if free_b < need:
            raise SystemExit(f"c5 needs {need >> 30} GiB on one GPU, only {free_b >> 30} GiB are free")
    sb.build()
    line = search_line(sb, args, nq, batches, with_cpu)
    sb.free()

    # ---- N = 1: the other BASELINE configs, measured in the same run (extra keys, own roofline / e2e / cpu) ----
    if world == 1 and name == "c5" and not (args.no_extras or args.no_scaling_base) and args.rows is None:
        sub = argparse.Namespace(**vars(args))
        sub.rows = None
        for other, bt in (("c2", [1, 64, 1024, 4096]), ("c3", [])):
            try:
                so = SearchBench(sub, other, torch, dist, bfk, rank, world, local_rank)
                so.build()
                line[other] = search_line(so, sub, SEARCH_WORKLOADS[other][5], bt, with_cpu, clocks_for_head=False)
                so.free()
            except Exception as e:  # an extra must not take the head line down with it
                line[other] = {"error": f"{type(e).__name__}: {e}"}
        ing = argparse.Namespace(**vars(sub))
        ing.rows = 300  # messages: enough for stable medians, a couple of seconds in all
        for other, fn in (("c4", lambda: run_c4(sub, with_cpu=with_cpu)), ("c4e", lambda: run_c4e(sub)), ("c1", lambda: run_c1(sub)),
                          ("c1i", lambda: run_c1i(ing))):
            try:
                line[other] = fn()
            except Exception as e:
                line[other] = {"error": f"{type(e).__name__}: {e}"}
            torch.cuda.empty_cache()

    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

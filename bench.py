#!/usr/bin/env python
"""Benchmark of the flat-kNN hot path (BASELINE.json metric: kNN queries/sec at k=10).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--nq Q] [--workload c2|c5] [--impl ours|reference]

N = 1 (default): BASELINE config 2 -- synthetic 1M x 384 fp32 unit-norm corpus, IndexFlatL2,
k = 10, one step = one search call over a batch of `nq` queries.  N > 1 (under torchrun,
one rank per GPU): BASELINE config 5 -- a fixed 100M x 384 bf16-stored corpus row-sharded
over the ranks (strong scaling), one all-gather of per-GPU top-k per step.

Prints ONE JSON line (rank 0).  `value` = queries/s with queries and results resident in
HBM; `e2e` = the same through the public numpy API with pinned host buffers (H2D + D2H
inside the timed region); `roofline` = algorithmic bytes of the scan kernel / its measured
launch time against MEASURED_PEAKS.json; `cpu_baseline` = the CPU oracle port on this box.
`--impl reference` times the CPU restatement of the reference path (faiss-cpu itself is not
installable here) on the same config.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

os.environ.setdefault("OMP_WAIT_POLICY", "passive")
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "kNN queries/sec at k=10"
K = 10
D = 384
METRIC_ID = 1  # faiss METRIC_L2; config 3 switches to inner product (0), d = 768, k = 100
CORPUS_SEED, QUERY_SEED = 1234, 4321


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), float(j.get("bf16_tflops", 1590.0)), "measured"
    return 6650.0, 1590.0, "fallback"


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


SPEC_HBM_GBS = 8000.0      # B200 data sheet; the north star quotes "~8 TB/s"
SPEC_BF16_TFLOPS = 2250.0  # dense bf16


def gen_unit_rows(torch, n, d, seed, device, chunk=1 << 20):
    """Seeded N(0, I) rows normalised to unit length, generated on the device in chunks."""
    g = torch.Generator(device=device).manual_seed(seed)
    for r0 in range(0, n, chunk):
        m = min(chunk, n - r0)
        x = torch.randn((m, d), generator=g, device=device, dtype=torch.float32)
        x /= x.norm(dim=1, keepdim=True)
        yield r0, x


def roofline_of(*a):
    """`_roofline_core` plus the fraction against the spec-sheet denominator (SURVEY 8d asks for both)."""
    r = _roofline_core(*a)
    r["frac_of_spec"] = r["achieved"] / (SPEC_HBM_GBS if r["bound"] == "hbm" else SPEC_BF16_TFLOPS)
    return r


def _roofline_core(res, nq, n_rows, elem, hbm_peak, tf_peak, peak_kind):
    """Roofline of the dominant (scan) kernel for one measured batch size.

    Algorithmic work per launch follows SURVEY 8(d) / DESIGN.md section 4:
      exact SIMT scan (<= 8 queries per corpus pass): N*d*s bytes + queries + per-CTA lists
      tensor scan, HBM regime (nq <= 128): N*d*2 (bf16 copy) + N*4 (norms) + nq*d*2 + lists
      tensor scan, tensor regime (nq > 128): 2*nq*N*d flop
    `launch_ms` is the library's own CUDA-event bracket around the scan launches of one search.
    """
    if res["algo"] == "exact_simt":
        passes = -(-nq // 8)
        per_q = min(nq, 8)
        bytes_launch = n_rows * D * elem + per_q * D * 4 + per_q * K * 8 * 148
        launch_ms = res["scan_ms"] / passes
        ach = bytes_launch / (launch_ms * 1e-3) / 1e9 if launch_ms > 0 else 0.0
        # profiles/r01_ncu_scan_exact_nq1.txt: dram read 1536.08 MB + write 3.75 MB at nq = 1
        traffic = 1539.83e6 if (n_rows == 1_000_000 and D == 384 and nq == 1 and elem == 4) else None
        return {"bound": "hbm", "kernel": "scan_exact_kernel", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                "frac": ach / hbm_peak, "peak_kind": peak_kind, "bytes_per_launch": bytes_launch,
                "launch_ms": launch_ms, "launches_per_step": passes, "traffic": traffic}
    launch_ms = res["scan_ms"]
    if nq <= 128:
        bytes_launch = n_rows * D * 2 + (n_rows * 4 if METRIC_ID == 1 else 0) + nq * D * 2 + nq * 148 * 16 * 12
        ach = bytes_launch / (launch_ms * 1e-3) / 1e9 if launch_ms > 0 else 0.0
        # DRAM bytes per launch from the one `ncu --set full` capture of this exact shape
        # (profiles/r01_ncu_scan_tc_final_nq64.txt: dram read 772.29 MB + write 4.24 MB)
        traffic = 776.53e6 if (n_rows == 1_000_000 and D == 384 and nq == 64 and METRIC_ID == 1) else None
        r = {"bound": "hbm", "kernel": "scan_tc_kernel", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
             "frac": ach / hbm_peak, "peak_kind": peak_kind, "bytes_per_launch": bytes_launch,
             "launch_ms": launch_ms, "launches_per_step": 1, "traffic": traffic}
        if r["frac"] > 1.0:
            # the measured peak is a COPY (half reads, half writes); a read-only stream can run above it
            r["note"] = "read-only scan above the measured copy (read+write) bandwidth"
        return r
    flops = 2.0 * nq * n_rows * D
    ach = flops / (launch_ms * 1e-3) / 1e12 if launch_ms > 0 else 0.0
    return {"bound": "tensor", "kernel": "scan_tc_kernel", "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s",
            "frac": ach / tf_peak, "peak_kind": peak_kind + " (burst bf16 GEMM)", "flops_per_launch": flops,
            "launch_ms": launch_ms, "launches_per_step": 1, "traffic": None}


def run_c4(args):
    """BASELINE config 4: encoder outputs [B, 256, 384] -> fused pool + normalise + cast -> index.add.

    One step = one add_pooled launch over a device batch of 16384 chunks; 4 rotating input
    buffers (3.2 GB each in bf16, so every launch streams from HBM, far above the 126 MB L2).
    """
    import torch

    import docqa_b200 as bfk

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    B, T, H = args.rows or 16384, 256, D
    hbm_peak, _, peak_kind = peaks()
    out = {}
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
            steps = max(args.steps, 8)
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
                                      "bytes_per_chunk": bytes_chunk, "achieved_gbs": ach, "frac": ach / hbm_peak}
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
                                      "achieved_gbs": ach, "frac": ach / hbm_peak, "kernel": "append_rows_kernel"}
        del index
    del xb
    torch.cuda.empty_cache()
    head = out["bf16_full"]
    line = {
        "metric": "ingestion chunks/sec (pool+normalise+cast+add)", "value": head["chunks_per_s"], "unit": "chunks/s",
        "n_gpus": 1, "steps": max(args.steps, 8), "warmup": args.warmup, "ms_per_step": head["ms_per_launch"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16 in, f32 accumulate", "data": "synthetic",
        "config": {"workload": f"c4: token states [{B}, {T}, {H}] bf16, int64 mask, fused pool+normalise+cast+index.add",
                   "l2_policy": "4 rotating 3.2 GB input buffers, far above the 126 MB L2"},
        "roofline": {"bound": "hbm", "kernel": "pool_norm_kernel", "achieved": head["achieved_gbs"], "peak": hbm_peak,
                     "unit": "GB/s", "frac": head["frac"], "peak_kind": peak_kind,
                     "bytes_per_launch": head["bytes_per_chunk"] * B, "launch_ms": head["ms_per_launch"], "traffic": None},
        "gpu_launches": max(args.steps, 8), "variants": out,
    }
    print(json.dumps(line))


def run_c1(args):
    """BASELINE config 1: the reference's own 649 x 384 index, one query per call, k = 5 (llm-qa/main.py:101,117).

    Too small for a roofline ('parity + latency only'): the number is the wall-clock latency of one
    `index.search(np.float32[1, 384], 5)` through the numpy API, host buffers in and out, next to the CPU
    restatement of faiss on the same box.
    """
    import docqa_b200 as bfk
    from oracle import cpu_baseline, faiss_io

    here = os.path.dirname(os.path.abspath(__file__))
    path = os.path.join(here, "tests", "golden", "docqa_vector_store.faiss")
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
    line = {
        "metric": "single-query search latency at k=5 (semantic-indexer default_data)", "value": float(np.median(lat)),
        "unit": "ms", "n_gpus": 1, "steps": 2000, "warmup": 50, "ms_per_step": float(np.median(lat)),
        "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "bf16 scan + f32 rerank",
        "data": "the reference's shipped vector_store.faiss (649 x 384)",
        "config": {"workload": "c1: 649x384 fixture, IndexFlatL2, k=5, nq=1 per call, numpy in / numpy out"},
        "latency_ms": {"p50": float(np.median(lat)), "p90": float(lat[int(0.9 * len(lat))]), "p99": float(lat[int(0.99 * len(lat))])},
        "e2e": {"value": float(np.median(lat)), "unit": "ms", "h2d_bytes_per_step": 384 * 4, "d2h_bytes_per_step": 5 * 12},
        "gpu_launches": 3 * 2000, "roofline": None,
        "cpu_baseline": {"value": float(np.median(cl)), "unit": "ms", "cores": 1, "kind": "port",
                         "sample": "300 single-query calls of the faiss restatement (seq path) on the same 649 rows"},
    }
    print(json.dumps(line))


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
    _, tf_peak, peak_kind = peaks()
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
    steps = max(args.steps, 5)
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
    line = {
        "metric": "ingestion chunks/sec (encoder + pool + normalise + add)", "value": B / (ms * 1e-3), "unit": "chunks/s",
        "n_gpus": 1, "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16 activations, f32 accumulate", "data": "synthetic",
        "config": {"workload": f"c4e: {B} chunks x {T} tokens, BERT 6x384 (12 heads, 1536) random weights -> mean pool -> index.add",
                   "l2_policy": "4 rotating id buffers; activations of a step (2 GB) far exceed the 126 MB L2"},
        "e2e": {"value": B / (ms_e2e * 1e-3), "unit": "chunks/s", "h2d_bytes_per_step": B * T * 16, "d2h_bytes_per_step": 8},
        "gpu_launches": steps * (1 + 5 * L + 1), "clocks": clocks,
        "roofline": {"bound": "tensor", "kernel": "enc_gemm_kernel (+ enc_attention_kernel)", "achieved": ach, "peak": tf_peak,
                     "unit": "TFLOP/s", "frac": ach / tf_peak, "frac_of_spec": ach / SPEC_BF16_TFLOPS,
                     "peak_kind": peak_kind + " (burst bf16 GEMM)", "flops_per_launch": (gemm_flop + attn_flop) * B,
                     "launch_ms": ms, "traffic": None,
                     "note": "whole forward pass (31 launches) against the tensor peak; per-kernel split in profiles/"},
        "cpu_baseline": None,
    }
    print(json.dumps(line))


def run_reference(args):
    """CPU arm: the oracle's restatement of faiss-cpu IndexFlatL2.search on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cpu_baseline

    cpu_baseline.build()
    workload = args.workload or ("c2" if args.gpus == 1 else "c5")
    n_full = {"c2": 1_000_000, "c3": 10_000_000}.get(workload, 100_000_000)
    n_sample = min(args.rows or 1_000_000, n_full)
    rng = np.random.default_rng(CORPUS_SEED)
    xb = rng.standard_normal((n_sample, D), dtype=np.float32)
    xb /= np.linalg.norm(xb, axis=1, keepdims=True)
    xq = np.random.default_rng(QUERY_SEED).standard_normal((args.nq, D), dtype=np.float32)
    xq /= np.linalg.norm(xq, axis=1, keepdims=True)
    bn = np.einsum("ij,ij->i", xb, xb, dtype=np.float32)

    # a real faiss-cpu is preferred the day one is importable (none is in this image)
    real = None
    try:
        import faiss as _faiss

        if getattr(_faiss, "__version__", "bfk-1.0") != "bfk-1.0" and hasattr(_faiss, "omp_get_max_threads"):
            real = _faiss.IndexFlatL2(D) if METRIC_ID == 1 else _faiss.IndexFlatIP(D)
            real.add(xb)
    except Exception:
        real = None

    def step():
        if real is not None:
            real.search(xq, K)
        elif args.nq < 20:
            cpu_baseline.knn_seq(xq, xb, K, METRIC_ID)
        else:
            cpu_baseline.knn_blas(xq, xb, K, METRIC_ID, bn=bn)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    scale = n_sample / n_full  # brute force is linear in N
    qps = args.nq * args.steps / dt * scale
    cores = min(args.nq, cpu_baseline.num_threads()) if args.nq < 20 else cpu_baseline.num_threads()
    kind, ref_kind = "port", "CPU restatement of faiss-cpu IndexFlatL2.search (faiss not installable here)"
    if real is not None:
        cores, kind, ref_kind = _faiss.omp_get_max_threads(), "reference", f"faiss-cpu {_faiss.__version__}"
    sample = (f"full workload: {n_sample}x{D} fp32 corpus, nq={args.nq}" if scale == 1.0 else
              f"{n_sample}-row slice of the {n_full}-row corpus, nq={args.nq}; value extrapolated linearly in N")
    line = {
        "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3 / scale, "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "impl": "reference",
        "config": {"workload": f"{workload}: {n_full}x{D} corpus, {'IndexFlatL2' if METRIC_ID == 1 else 'IndexFlatIP'}, k={K}, nq={args.nq}",
                   "reference_kind": ref_kind},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--nq", type=int, default=None)
    ap.add_argument("--workload", choices=["c1", "c2", "c3", "c5", "c4", "c4e"], default=None)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--rows", type=int, default=None, help="override corpus rows (debugging)")
    ap.add_argument("--algo", default=None, choices=[None, "auto", "exact", "tensor"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-scaling-base", action="store_true", help="skip the one-GPU run of config 5 at N=1")
    ap.add_argument("--batches", default="1,64,1024,4096", help="extra batch sizes reported under `batches`")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    global D, K, METRIC_ID, METRIC
    if args.workload == "c3":  # BASELINE config 3: 10M x 768 bf16-stored, IndexFlatIP, batch 4096, k = 100
        D, K, METRIC_ID, METRIC = 768, 100, 0, "kNN queries/sec at k=100"
        args.nq = args.nq or 4096
        if args.batches == "1,64,1024,4096":
            args.batches = ""
    args.nq = args.nq or 64
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "c1":
        return run_c1(args)
    if args.workload == "c4":
        return run_c4(args)
    if args.workload == "c4e":
        return run_c4e(args)

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

    workload = args.workload or ("c2" if world == 1 else "c5")
    if workload == "c5" and args.batches == "1,64,1024,4096":
        args.batches = "1,64,1024,8192"  # C5 batches
    if workload == "c2":
        n_total, storage, scaling = args.rows or 1_000_000, "f32", "weak" if world == 1 else "strong"
    elif workload == "c3":
        n_total, storage, scaling = args.rows or 10_000_000, "bf16", "weak" if world == 1 else "strong"
    else:
        n_total, storage, scaling = args.rows or 100_000_000, "bf16", "strong"
    lo, hi = bfk.shard_range(n_total, rank, world)

    # ---- build the (sharded) index on the device ----
    if world > 1:
        index = bfk.ShardedIndexFlat(D, METRIC_ID, device=local_rank, storage=storage)
        local = index.local
    else:
        index = local = bfk.IndexFlat(D, METRIC_ID, device=local_rank, storage=storage)
    local.reserve(hi - lo)
    # every rank draws the same global stream and keeps its slice, so the corpus does not depend on N
    for r0, x in gen_unit_rows(torch, n_total, D, CORPUS_SEED, dev):
        a, b = max(lo, r0), min(hi, r0 + x.shape[0])
        if b > a:
            local.add(x[a - r0 : b - r0])
        del x
    if world > 1:
        index._record(hi - lo, lo, n_total)
    torch.cuda.synchronize()
    assert local.ntotal == hi - lo

    def queries(nq):
        g = torch.Generator(device=dev).manual_seed(QUERY_SEED)
        q = torch.randn((nq, D), generator=g, device=dev, dtype=torch.float32)
        return q / q.norm(dim=1, keepdim=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def time_steps(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    def measure(nq, steps, warmup, with_e2e=True, sustain_s=0.0):
        q_dev = queries(nq)
        D_dev = torch.empty((nq, K), dtype=torch.float32, device=dev)
        I_dev = torch.empty((nq, K), dtype=torch.int64, device=dev)
        kw = {"algo": args.algo} if args.algo else {}

        if world == 1:
            def dev_step():
                index.search(q_dev, K, D=D_dev, I=I_dev, **kw)
        else:
            def dev_step():
                index.search(q_dev, K, **kw)

        ms = time_steps(dev_step, steps, warmup)
        out = {"nq": nq, "value": nq * steps / (ms * 1e-3), "ms_per_step": ms / steps}
        if sustain_s > 0:
            # the K timed steps last only milliseconds; keep the same step running long enough for
            # nvidia-smi (100 ms period) to see the clocks the kernel actually runs at, and report
            # the throughput of that longer window too
            n_rep = max(steps, int(sustain_s / max(ms / steps * 1e-3, 1e-6)))
            ms_s = time_steps(dev_step, n_rep, 0)
            out["sustained"] = {"value": nq * n_rep / (ms_s * 1e-3), "steps": n_rep, "seconds": ms_s * 1e-3}
        # launch count + scan-kernel time from the library's own events (local shard)
        st = bfk.SearchStats()
        scan_ms, launches = [], 0
        for _ in range(min(steps, 10)):
            local.search(q_dev, K, stats=st, **kw)
            scan_ms.append(st.scan_ms)
            launches = st.n_launches
        out["scan_ms"] = statistics.mean(scan_ms)
        out["launches_per_step"] = launches + (4 if world > 1 else 0)  # + K6x push, wait, pack, select
        out["algo"] = {1: "exact_simt", 2: "tensor"}.get(st.algo_used, "?")
        out["n_fallback"] = int(st.n_fallback)
        if with_e2e:
            q_pin = torch.empty((nq, D), dtype=torch.float32, pin_memory=True)
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

            ms_h = time_steps(host_step, steps, warmup)
            out["e2e"] = {"value": nq * steps / (ms_h * 1e-3), "unit": "queries/s",
                          "h2d_bytes_per_step": nq * D * 4, "d2h_bytes_per_step": nq * K * 12}
        return out

    sampler = ClockSampler(local_rank) if rank == 0 else None
    time.sleep(0.3)  # let nvidia-smi start sampling
    head = measure(args.nq, args.steps, args.warmup, sustain_s=1.5)
    clocks = sampler.stop() if sampler else None
    extra = {}
    for b in [int(s) for s in args.batches.split(",") if s]:
        if b != args.nq:
            r = measure(b, max(3, min(args.steps, 10)), 3, with_e2e=True)
            extra[str(b)] = {"value": r["value"], "ms_per_step": r["ms_per_step"], "algo": r["algo"],
                             "e2e": r["e2e"]["value"], "scan_ms": r["scan_ms"]}

    # ---- roofline of the dominant (scan) kernel ----
    hbm_peak, tf_peak, peak_kind = peaks()
    elem = 4 if storage == "f32" else 2
    n_local = hi - lo
    roofline = roofline_of(head, args.nq, n_local, elem, hbm_peak, tf_peak, peak_kind)
    for b, r in extra.items():
        r["roofline"] = roofline_of({"algo": r["algo"], "scan_ms": r["scan_ms"]}, int(b), n_local, elem, hbm_peak,
                                    tf_peak, peak_kind)

    # ---- CPU baseline on this box (rank 0, N = 1 only) ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cpu_baseline

        cpu_baseline.build()
        xb_h = local.reconstruct_n(0, min(n_local, 1_000_000))
        xq_h = queries(args.nq).cpu().numpy()
        bn = np.einsum("ij,ij->i", xb_h, xb_h, dtype=np.float32)
        fn = (lambda: cpu_baseline.knn_seq(xq_h, xb_h, K, METRIC_ID)) if args.nq < 20 else (
            lambda: cpu_baseline.knn_blas(xq_h, xb_h, K, METRIC_ID, bn=bn))
        fn()
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 10.0 and reps < 50:
            fn()
            reps += 1
        dt = time.perf_counter() - t0
        cores = min(args.nq, cpu_baseline.num_threads()) if args.nq < 20 else cpu_baseline.num_threads()
        cpu = {"value": args.nq * reps / dt * (xb_h.shape[0] / n_total), "unit": "queries/s", "cores": cores,
               "kind": "port", "host_cpus": os.cpu_count(),
               "sample": f"{reps} calls of nq={args.nq} over {xb_h.shape[0]}x{D} fp32 rows "
                         f"({'seq' if args.nq < 20 else 'blas'} path of the faiss restatement)"}

    # ---- same-workload base for the multi-GPU lines (N = 1 only) ----
    # `--gpus N` for N > 1 runs config 5 (100M x 384 bf16-stored, strong scaling); its one-GPU point is
    # measured here, on the same corpus stream, so that value(N) / c5_single_gpu.value is a like-for-like
    # scaling figure (the c2 head line above is a different, 100x smaller corpus).
    c5_base = None
    if world == 1 and workload == "c2" and not args.no_scaling_base and args.rows is None:
        free_b, _ = torch.cuda.mem_get_info(dev)
        need = 100_000_000 * (D * 2 + 4) + (8 << 30)
        if free_b > need:
            del index, local
            torch.cuda.empty_cache()
            n_total5 = 100_000_000
            index = local = bfk.IndexFlat(D, METRIC_ID, device=local_rank, storage="bf16")
            local.reserve(n_total5)
            for r0, x in gen_unit_rows(torch, n_total5, D, CORPUS_SEED, dev):
                local.add(x)
                del x
            torch.cuda.synchronize()
            c5_base = {"workload": f"c5 on one GPU: {n_total5}x{D} bf16-stored corpus, IndexFlatL2, k={K}", "batches": {}}
            for b in (64, 1, 1024):
                r = measure(b, 10 if b <= 64 else 3, 3, with_e2e=(b == 64))
                c5_base["batches"][str(b)] = {
                    "value": r["value"], "ms_per_step": r["ms_per_step"], "scan_ms": r["scan_ms"], "algo": r["algo"],
                    "roofline": roofline_of({"algo": r["algo"], "scan_ms": r["scan_ms"]}, b, n_total5, 2, hbm_peak,
                                            tf_peak, peak_kind)}
                if b == 64:
                    c5_base["value"], c5_base["nq"], c5_base["e2e"] = r["value"], 64, r["e2e"]["value"]
        else:
            c5_base = {"skipped": f"only {free_b >> 30} GiB free on the device"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": head["value"], "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f32" if head["algo"] == "exact_simt" else "bf16 scan + f32 rerank",
            "data": "synthetic",
            "config": {"workload": f"{workload}: {n_total}x{D} {'fp32' if storage == 'f32' else 'bf16-stored'} corpus, "
                                   f"{'IndexFlatL2' if METRIC_ID == 1 else 'IndexFlatIP'}, k={K}, nq={args.nq}" + (f", row-sharded over {world} GPUs" if world > 1 else ""),
                       "l2_policy": "corpus pass (>= 0.77 GB/GPU at N=1) far exceeds the 126 MB L2; no flush needed",
                       "algo": head["algo"]},
            "e2e": head["e2e"], "gpu_launches": head["launches_per_step"] * args.steps, "clocks": clocks,
            "roofline": roofline, "cpu_baseline": cpu, "batches": extra, "n_fallback": head["n_fallback"],
            "sustained": head.get("sustained"),
        }
        if c5_base is not None:
            line["c5_single_gpu"] = c5_base
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

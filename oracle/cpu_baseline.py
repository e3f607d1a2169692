"""ORACLE (test infrastructure, not product code): threaded CPU flat kNN.

Drives oracle/libflat_ref.so (C, OpenMP) the way faiss-cpu dispatches
`IndexFlat::search` (reached from llm-qa/main.py:101,117): below 20 queries the
sequential direct-distance scan, one OpenMP iteration per query; from 20 queries
the blocked BLAS form (4096 queries x 1024 rows per sgemm, numpy/OpenBLAS doing
the sgemm, the C heap handler folding each block).  This is bench.py's
`cpu_baseline` (kind "port") and its `--impl reference` arm, and a fast checker
for tests; it is PARITY UNPINNED against live faiss like the rest of oracle/.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libflat_ref.so")
    src = os.path.join(_HERE, "flat_ref.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "libflat_ref.so"])
    return so


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libflat_ref.so")
        if not os.path.exists(so):
            build()
        L = ctypes.CDLL(so)
        f32p = ctypes.POINTER(ctypes.c_float)
        i64p = ctypes.POINTER(ctypes.c_int64)
        i64 = ctypes.c_int64
        L.flatref_knn_seq.argtypes = [f32p, i64, f32p, i64, ctypes.c_int, i64, ctypes.c_int, f32p, i64p]
        L.flatref_block_update.argtypes = [f32p, i64, i64, i64, i64, i64, ctypes.c_int, f32p, f32p, f32p, i64p]
        L.flatref_heap_init.argtypes = [i64, i64, f32p, i64p]
        L.flatref_heap_finish.argtypes = [i64, i64, ctypes.c_int, f32p, i64p]
        L.flatref_pool_normalize.argtypes = [f32p, i64p, i64, ctypes.c_int, ctypes.c_int, f32p]
        L.flatref_num_threads.restype = ctypes.c_int
        _LIB = L
    return _LIB


def _f(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _i(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))


def num_threads() -> int:
    return int(lib().flatref_num_threads())


def knn_seq(xq, xb, k, metric=METRIC_L2):
    xq = np.ascontiguousarray(xq, dtype=np.float32)
    xb = np.ascontiguousarray(xb, dtype=np.float32)
    nq, d = xq.shape
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    rc = lib().flatref_knn_seq(_f(xq), nq, _f(xb), xb.shape[0], d, k, metric, _f(D), _i(I))
    if rc:
        raise RuntimeError("flatref_knn_seq failed")
    return D, I


def knn_blas(xq, xb, k, metric=METRIC_L2, query_bs=4096, database_bs=1024, bn=None):
    xq = np.ascontiguousarray(xq, dtype=np.float32)
    xb = np.ascontiguousarray(xb, dtype=np.float32)
    nq, d = xq.shape
    n = xb.shape[0]
    L = lib()
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    L.flatref_heap_init(nq, k, _f(D), _i(I))
    qn = np.einsum("ij,ij->i", xq, xq, dtype=np.float32)
    if bn is None:
        bn = np.einsum("ij,ij->i", xb, xb, dtype=np.float32)
    for i0 in range(0, nq, query_bs):
        i1 = min(nq, i0 + query_bs)
        Dv = D[i0:i1]
        Iv = I[i0:i1]
        for j0 in range(0, n, database_bs):
            j1 = min(n, j0 + database_bs)
            ip = xq[i0:i1] @ xb[j0:j1].T
            L.flatref_block_update(_f(ip), i1 - i0, j1 - j0, ip.shape[1], j0, k, metric,
                                   _f(qn[i0:i1]), _f(bn[j0:j1]), _f(Dv), _i(Iv))
    L.flatref_heap_finish(nq, k, metric, _f(D), _i(I))
    return D, I


def knn(xq, xb, k, metric=METRIC_L2):
    """faiss dispatch (distance_compute_blas_threshold = 20)."""
    if np.asarray(xq).shape[0] < 20:
        return knn_seq(xq, xb, k, metric)
    return knn_blas(xq, xb, k, metric)


def pool_normalize(tok, mask):
    tok = np.ascontiguousarray(tok, dtype=np.float32)
    mask = np.ascontiguousarray(mask, dtype=np.int64)
    B, T, H = tok.shape
    out = np.empty((B, H), np.float32)
    lib().flatref_pool_normalize(_f(tok), _i(mask), B, T, H, _f(out))
    return out

"""ORACLE (test infrastructure, not product code): exact flat kNN on the CPU.

PARITY UNPINNED against live faiss: the arithmetic of this path lives in the
third-party `faiss-cpu` wheel (unpinned, `semantic-indexer/requirements.txt:2`,
`llm-qa/requirements.txt:6`), which is neither vendored under /root/reference
nor installable in this image, and the reference's own test mocks it wholesale
(`semantic-indexer/tests/test_indexer.py:19-20`).  This file restates faiss's
published IndexFlat algorithm and is pinned only to what the reference ships:
the `vector_store.faiss` artefact (tests/golden/) and its call sites
    indexer.py:39,104   faiss.IndexFlatL2(dimension)
    indexer.py:41       index.add(float32[n, d])
    llm-qa/main.py:101  index.search(float32[1, d], k=3) via LangChain
    indexer.py:109      index.ntotal

Restated behaviour (faiss >= 1.7.3, IndexFlat::search -> knn_L2sqr /
knn_inner_product):
  * L2 returns SQUARED distances, best (smallest) first; IP returns inner
    products, best (largest) first.
  * nq < 20  ("seq" path, exhaustive_L2sqr_seq): direct sum((x - y)^2) in fp32.
  * nq >= 20 ("blas" path, exhaustive_L2sqr_blas): ||x||^2 + ||y||^2 - 2<x,y>
    from an sgemm over blocks of 4096 queries x 1024 rows, negatives clamped
    to 0.
  * selection is exact; a candidate enters only if strictly better than the
    current k-th; final lists are sorted; equal distances are ordered by id.
  * missing results (k > ntotal): I = -1, D = +FLT_MAX (L2) / -FLT_MAX (IP).
  * I is int64, D float32, both [nq, k] C-contiguous.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may
import this module.
"""
from __future__ import annotations

import numpy as np

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1
FLT_MAX = np.float32(3.4028234663852886e38)

# faiss globals: distance_compute_blas_threshold / _query_bs / _database_bs
BLAS_THRESHOLD = 20
BLAS_QUERY_BS = 4096
BLAS_DATABASE_BS = 1024


def _topk_rows(keys: np.ndarray, ids: np.ndarray, k: int):
    """Per row: the k smallest keys ordered by (key, id). keys [nq, m]."""
    nq, m = keys.shape
    kk = min(k, m)
    if kk < m:
        # argpartition is not tie-stable: widen the cut to every key <= k-th
        part = np.partition(keys, kk - 1, axis=1)[:, kk - 1 : kk]
        out_k = np.empty((nq, kk), dtype=keys.dtype)
        out_i = np.empty((nq, kk), dtype=np.int64)
        for r in range(nq):
            sel = np.nonzero(keys[r] <= part[r, 0])[0]
            order = np.lexsort((ids[r, sel] if ids.ndim == 2 else ids[sel], keys[r, sel]))[:kk]
            out_k[r] = keys[r, sel[order]]
            out_i[r] = (ids[r, sel[order]] if ids.ndim == 2 else ids[sel[order]])
        return out_k, out_i
    out_k = np.empty((nq, kk), dtype=keys.dtype)
    out_i = np.empty((nq, kk), dtype=np.int64)
    for r in range(nq):
        rid = ids[r] if ids.ndim == 2 else ids
        order = np.lexsort((rid, keys[r]))
        out_k[r] = keys[r, order]
        out_i[r] = rid[order]
    return out_k, out_i


def _pad(D, I, k, metric):
    nq, kk = D.shape
    if kk == k:
        return np.ascontiguousarray(D, dtype=np.float32), np.ascontiguousarray(I, dtype=np.int64)
    Do = np.full((nq, k), FLT_MAX if metric == METRIC_L2 else -FLT_MAX, dtype=np.float32)
    Io = np.full((nq, k), -1, dtype=np.int64)
    Do[:, :kk] = D
    Io[:, :kk] = I
    return Do, Io


def pairwise_seq(xq: np.ndarray, xb: np.ndarray, metric: int) -> np.ndarray:
    """fp32 distances the way the nq<20 path forms them (direct difference)."""
    xq = np.asarray(xq, dtype=np.float32)
    xb = np.asarray(xb, dtype=np.float32)
    out = np.empty((xq.shape[0], xb.shape[0]), dtype=np.float32)
    for i in range(xq.shape[0]):
        if metric == METRIC_L2:
            diff = xb - xq[i]
            out[i] = np.einsum("ij,ij->i", diff, diff, dtype=np.float32)
        else:
            out[i] = xb @ xq[i]
    return out


def pairwise_f64(xq: np.ndarray, xb: np.ndarray, metric: int) -> np.ndarray:
    """Reference distances in float64 (the 'truth' tolerance checks use)."""
    xq = np.asarray(xq, dtype=np.float64)
    xb = np.asarray(xb, dtype=np.float64)
    if metric == METRIC_L2:
        qn = (xq * xq).sum(1)[:, None]
        bn = (xb * xb).sum(1)[None, :]
        if xq.shape[0] * xb.shape[0] * xb.shape[1] <= 1 << 26:
            diff = xq[:, None, :] - xb[None, :, :]
            return (diff * diff).sum(-1)
        return np.maximum(qn + bn - 2.0 * (xq @ xb.T), 0.0)
    return xq @ xb.T


def knn_seq(xq, xb, k, metric=METRIC_L2):
    """nq < 20 path: direct distances + exact selection."""
    xq = np.ascontiguousarray(xq, dtype=np.float32)
    xb = np.ascontiguousarray(xb, dtype=np.float32)
    nq, n = xq.shape[0], xb.shape[0]
    if n == 0 or nq == 0:
        return _pad(np.empty((nq, 0), np.float32), np.empty((nq, 0), np.int64), k, metric)
    dis = pairwise_seq(xq, xb, metric)
    keys = dis if metric == METRIC_L2 else -dis
    K, I = _topk_rows(keys, np.arange(n, dtype=np.int64), k)
    D = K if metric == METRIC_L2 else -K
    return _pad(D.astype(np.float32), I, k, metric)


def knn_blas(xq, xb, k, metric=METRIC_L2, query_bs=BLAS_QUERY_BS, database_bs=BLAS_DATABASE_BS):
    """nq >= 20 path: blocked sgemm + norms, clamp, exact selection."""
    xq = np.ascontiguousarray(xq, dtype=np.float32)
    xb = np.ascontiguousarray(xb, dtype=np.float32)
    nq, n = xq.shape[0], xb.shape[0]
    if n == 0 or nq == 0:
        return _pad(np.empty((nq, 0), np.float32), np.empty((nq, 0), np.int64), k, metric)
    qn = np.einsum("ij,ij->i", xq, xq, dtype=np.float32)
    bn = np.einsum("ij,ij->i", xb, xb, dtype=np.float32)
    Dall = np.empty((nq, 0), np.float32)
    Iall = np.empty((nq, 0), np.int64)
    outD, outI = [], []
    for i0 in range(0, nq, query_bs):
        i1 = min(nq, i0 + query_bs)
        bestK = np.empty((i1 - i0, 0), np.float32)
        bestI = np.empty((i1 - i0, 0), np.int64)
        for j0 in range(0, n, database_bs):
            j1 = min(n, j0 + database_bs)
            ip = xq[i0:i1] @ xb[j0:j1].T
            if metric == METRIC_L2:
                dis = qn[i0:i1, None] + bn[None, j0:j1] - np.float32(2.0) * ip
                np.maximum(dis, np.float32(0.0), out=dis)
                keys = dis
            else:
                keys = -ip
            ids = np.broadcast_to(np.arange(j0, j1, dtype=np.int64), keys.shape)
            catK = np.concatenate([bestK, keys], axis=1)
            catI = np.concatenate([bestI, ids], axis=1)
            bestK, bestI = _topk_rows(catK, catI, k)
        outD.append(bestK if metric == METRIC_L2 else -bestK)
        outI.append(bestI)
    Dall = np.concatenate(outD, axis=0)
    Iall = np.concatenate(outI, axis=0)
    return _pad(Dall.astype(np.float32), Iall, k, metric)


def knn(xq, xb, k, metric=METRIC_L2):
    """faiss dispatch: seq below the BLAS threshold, blas at or above it."""
    assert k > 0
    xq = np.asarray(xq)
    if xq.shape[0] < BLAS_THRESHOLD:
        return knn_seq(xq, xb, k, metric)
    return knn_blas(xq, xb, k, metric)


def knn_f64(xq, xb, k, metric=METRIC_L2):
    """Exact top-k by float64 distances, ties by id. Returns (D64, I)."""
    xb = np.asarray(xb)
    nq, n = np.asarray(xq).shape[0], xb.shape[0]
    if n == 0 or nq == 0:
        return (np.full((nq, k), np.inf if metric == METRIC_L2 else -np.inf),
                np.full((nq, k), -1, np.int64))
    dis = pairwise_f64(xq, xb, metric)
    keys = dis if metric == METRIC_L2 else -dis
    K, I = _topk_rows(keys, np.arange(n, dtype=np.int64), k)
    D = K if metric == METRIC_L2 else -K
    if D.shape[1] < k:
        Dp = np.full((nq, k), np.inf if metric == METRIC_L2 else -np.inf)
        Ip = np.full((nq, k), -1, np.int64)
        Dp[:, : D.shape[1]] = D
        Ip[:, : D.shape[1]] = I
        return Dp, Ip
    return D, I


def bitmap_to_mask(bitmap: np.ndarray, n: int) -> np.ndarray:
    """faiss IDSelectorBitmap layout: row i admitted <=> bit (i & 7) of byte (i >> 3) is set."""
    bits = np.unpackbits(np.ascontiguousarray(bitmap, dtype=np.uint8), bitorder="little")
    out = np.zeros(n, dtype=bool)
    m = min(n, bits.size)
    out[:m] = bits[:m].astype(bool)
    return out


def knn_selected(xq, xb, k, metric, mask):
    """faiss search with SearchParameters(sel=...): the rows the selector rejects do not exist for the search
    (IndexFlat::search hands `sel` to the distance computer loop; ids stay the original row numbers)."""
    ids = np.flatnonzero(mask)
    if ids.size == 0:
        D = np.empty((xq.shape[0], 0), dtype=np.float32)
        I = np.empty((xq.shape[0], 0), dtype=np.int64)
        return _pad(D, I, k, metric)
    D, I = knn(xq, np.ascontiguousarray(xb[ids]), k, metric)
    I = np.where(I >= 0, ids[np.clip(I, 0, ids.size - 1)], -1).astype(np.int64)
    return D, I


def range_search(xq, xb, radius, metric):
    """faiss IndexFlat::range_search: squared L2 < radius / inner product > radius (strict), direct fp32
    arithmetic (RangeSearchResultHandler over exhaustive_L2sqr_seq / exhaustive_inner_product_seq).
    Returns (lims int64[nq+1], D float32[total], I int64[total]); rows of a query in id order."""
    nq = xq.shape[0]
    lims = np.zeros(nq + 1, dtype=np.int64)
    Ds, Is = [], []
    for i in range(nq):
        dist = pairwise_seq(xq[i : i + 1], xb, metric)[0] if xb.shape[0] else np.zeros(0, dtype=np.float32)
        keep = dist < np.float32(radius) if metric == METRIC_L2 else dist > np.float32(radius)
        ids = np.flatnonzero(keep).astype(np.int64)
        Ds.append(dist[ids].astype(np.float32))
        Is.append(ids)
        lims[i + 1] = lims[i] + ids.size
    D = np.concatenate(Ds) if Ds else np.zeros(0, dtype=np.float32)
    I = np.concatenate(Is) if Is else np.zeros(0, dtype=np.int64)
    return lims, D, I


class IndexFlatRef:
    """CPU mirror of the faiss.IndexFlat surface the reference touches.

    Follows `semantic-indexer/indexer.py:32-48` (float32 cast, sequential ids)
    and the LangChain side paths listed in SURVEY.md section 8b.
    """

    def __init__(self, d: int, metric: int = METRIC_L2):
        self.d = int(d)
        self.metric_type = int(metric)
        self.is_trained = True
        self._x = np.empty((0, self.d), dtype=np.float32)

    @property
    def ntotal(self) -> int:
        return self._x.shape[0]

    def add(self, x):
        x = np.asarray(x)
        n, d = x.shape
        assert d == self.d
        self._x = np.concatenate([self._x, np.ascontiguousarray(x, dtype=np.float32)], axis=0)

    def search(self, x, k, *, params=None):
        x = np.asarray(x)
        n, d = x.shape
        assert d == self.d
        assert k > 0
        sel = getattr(params, "sel", None) if params is not None else None
        if sel is not None:
            return knn_selected(np.ascontiguousarray(x, dtype=np.float32), self._x, k, self.metric_type,
                                bitmap_to_mask(sel.to_bitmap(self.ntotal), self.ntotal))
        return knn(np.ascontiguousarray(x, dtype=np.float32), self._x, k, self.metric_type)

    def range_search(self, x, thresh):
        return range_search(np.ascontiguousarray(np.asarray(x), dtype=np.float32), self._x, float(thresh),
                            self.metric_type)

    def reset(self):
        self._x = np.empty((0, self.d), dtype=np.float32)

    def reconstruct(self, i):
        if not 0 <= i < self.ntotal:
            raise RuntimeError("reconstruct: id out of range")
        return self._x[i].copy()

    def reconstruct_n(self, i0, n):
        if not (0 <= i0 and i0 + n <= self.ntotal):
            raise RuntimeError("reconstruct_n: range out of bounds")
        return self._x[i0 : i0 + n].copy()

    def remove_ids(self, ids) -> int:
        ids = np.unique(np.asarray(ids, dtype=np.int64))
        ids = ids[(ids >= 0) & (ids < self.ntotal)]
        keep = np.ones(self.ntotal, dtype=bool)
        keep[ids] = False
        self._x = np.ascontiguousarray(self._x[keep])
        return int(ids.size)

    def merge_from(self, other):
        assert other.d == self.d
        self.add(other._x)
        other.reset()


def normalize_L2(x: np.ndarray) -> None:
    """faiss.normalize_L2 (fvec_renorm_L2): in place; zero rows are left alone."""
    assert x.dtype == np.float32 and x.flags.c_contiguous
    for i in range(x.shape[0]):
        nr = np.float32(np.dot(x[i], x[i]))
        if nr > 0:
            x[i] *= np.float32(1.0) / np.sqrt(nr, dtype=np.float32)

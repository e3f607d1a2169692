"""ORACLE (test infrastructure, not product code): masked mean-pool + L2 normalise.

PARITY UNPINNED against the live library: this is the tail of
`SentenceTransformer('all-MiniLM-L6-v2').encode(...)` as reached from
`semantic-indexer/indexer.py:21,37` and `llm-qa/main.py:25`; the
`sentence-transformers` wheel is unpinned, absent from /root/reference and not
installable here, so its published Pooling(mean) + Normalize modules are
restated:

    e = sum_t tok[t] * m[t] / max(sum_t m[t], 1e-9)       (models.Pooling, mean)
    v = e / max(||e||_2, 1e-12)                            (models.Normalize)

The only numeric pin the reference gives is indirect: every row of the shipped
`vector_store.faiss` is unit-norm to 1e-7 (tests/test_oracle_pool.py).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may
import this module.
"""
from __future__ import annotations

import numpy as np


def mean_pool_normalize(tok: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """tok [B, T, H] (any float dtype), mask [B, T] (0/1) -> [B, H] float32."""
    tok = np.asarray(tok).astype(np.float32)
    m = np.asarray(mask).astype(np.float32)
    s = np.einsum("bth,bt->bh", tok, m, dtype=np.float32)
    cnt = np.maximum(m.sum(axis=1, dtype=np.float32), np.float32(1e-9))
    e = s / cnt[:, None]
    nrm = np.sqrt(np.einsum("bh,bh->b", e, e, dtype=np.float32))
    return (e / np.maximum(nrm, np.float32(1e-12))[:, None]).astype(np.float32)


def mean_pool_normalize_f64(tok: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """Same formulas in float64, for tolerance checks."""
    tok = np.asarray(tok).astype(np.float64)
    m = np.asarray(mask).astype(np.float64)
    s = np.einsum("bth,bt->bh", tok, m)
    cnt = np.maximum(m.sum(axis=1), 1e-9)
    e = s / cnt[:, None]
    nrm = np.sqrt((e * e).sum(1))
    return e / np.maximum(nrm, 1e-12)[:, None]

"""Parity comparator for (D, I) results against float64-exact distances.

The bar (BASELINE.json north_star): ids identical to the reference except inside distance
ties within 1e-4 relative; distances within 1e-3 relative.  Near zero a relative bound is
meaningless (duplicates give 0 or ~1e-7 depending on faiss's code path), so distances get
the absolute floor of SURVEY.md section 7: 4*eps32*(||q||^2 + ||x||^2).
"""
from __future__ import annotations

import numpy as np

from oracle import flat_ref

RTOL_D = 1e-3
RTOL_TIE = 1e-4
EPS32 = float(np.finfo(np.float32).eps)
FLT_MAX = float(np.finfo(np.float32).max)


def check_knn(D, I, xq, xb, k, metric, *, I_ref=None, D_ref64=None, tight=False):
    """Assert (D, I) is a valid exact top-k answer for (xq, xb).

    I_ref / D_ref64: optional precomputed float64 top-k (flat_ref.knn_f64) to compare ids
    position by position.  `tight` additionally demands fp32-rounding-level distance
    agreement (for the direct-difference path).
    """
    xq = np.asarray(xq, dtype=np.float64)
    xb = np.asarray(xb, dtype=np.float64)
    nq, n = xq.shape[0], xb.shape[0]
    assert D.shape == (nq, k) and I.shape == (nq, k)
    assert D.dtype == np.float32 and I.dtype == np.int64
    kk = min(k, n)
    if I_ref is None and n > 0:
        D_ref64, I_ref = flat_ref.knn_f64(xq, xb, k, metric)
    qn = (xq * xq).sum(1)
    bn = (xb * xb).sum(1) if n else np.zeros(0)
    pad_d = FLT_MAX if metric == flat_ref.METRIC_L2 else -FLT_MAX
    for r in range(nq):
        ids = I[r, :kk]
        assert (I[r, kk:] == -1).all(), f"row {r}: padding ids must be -1"
        assert (D[r, kk:] == np.float32(pad_d)).all(), f"row {r}: padding distances must be +/-FLT_MAX"
        if kk == 0:
            continue
        assert ids.min() >= 0 and ids.max() < n, f"row {r}: id out of range"
        assert len(set(ids.tolist())) == kk, f"row {r}: duplicate ids {ids}"
        if metric == flat_ref.METRIC_L2:
            true = ((xb[ids] - xq[r]) ** 2).sum(1)
        else:
            true = xb[ids] @ xq[r]
        scale = qn[r] + bn[ids]
        atol = (4 * EPS32 if tight else 64 * EPS32) * np.maximum(scale, 1e-30)
        rtol = 8 * EPS32 if tight else RTOL_D
        err = np.abs(D[r, :kk].astype(np.float64) - true)
        assert (err <= rtol * np.abs(true) + atol).all(), (
            f"row {r}: distance mismatch max err {err.max():.3e} (true {true}, got {D[r, :kk]})")
        key = true if metric == flat_ref.METRIC_L2 else -true
        tie_tol = RTOL_TIE * np.abs(key) + 64 * EPS32 * scale
        # best-first order, up to ties
        assert (np.diff(key) >= -(tie_tol[1:] + tie_tol[:-1])).all(), f"row {r}: not sorted best-first: {true}"
        # position-wise id agreement except inside ties
        ref_key = D_ref64[r, :kk] if metric == flat_ref.METRIC_L2 else -D_ref64[r, :kk]
        diff = ids != I_ref[r, :kk]
        if diff.any():
            gap = np.abs(key - ref_key)
            assert (gap[diff] <= tie_tol[diff]).all(), (
                f"row {r}: ids differ outside a distance tie: got {ids}, want {I_ref[r, :kk]}, gaps {gap}")
    return True

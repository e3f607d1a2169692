"""Property tests (hypothesis) of the CPU oracle and the selector algebra: CPU only, small shapes."""
import numpy as np
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import faiss_io, flat_ref

L2, IP = flat_ref.METRIC_L2, flat_ref.METRIC_INNER_PRODUCT
FLT_MAX = np.finfo(np.float32).max


def _arr(seed, n, d):
    return np.random.default_rng(seed).standard_normal((n, d)).astype(np.float32)


@settings(max_examples=40, deadline=None)
@given(seed=st.integers(0, 10_000), n=st.integers(0, 60), d=st.integers(1, 24), nq=st.integers(1, 25), k=st.integers(1, 12),
       metric=st.sampled_from([L2, IP]))
def test_knn_shape_order_padding_and_agreement_of_both_paths(seed, n, d, nq, k, metric):
    xb, xq = _arr(seed, n, d), _arr(seed + 1, nq, d)
    D, I = flat_ref.knn(xq, xb, k, metric)
    assert D.shape == (nq, k) and I.shape == (nq, k) and D.dtype == np.float32 and I.dtype == np.int64
    kk = min(k, n)
    assert (I[:, kk:] == -1).all() and (D[:, kk:] == (FLT_MAX if metric == L2 else -FLT_MAX)).all()
    if n:
        key = D[:, :kk] if metric == L2 else -D[:, :kk]
        assert (np.diff(key, axis=1) >= 0).all()                       # best first
        assert all(len(set(r[:kk])) == kk for r in I.tolist())          # no row twice
        assert ((I[:, :kk] >= 0) & (I[:, :kk] < n)).all()
        # the sequential (nq < 20) and the BLAS (nq >= 20) forms return the same neighbours up to rounding ties
        Ds, Is = flat_ref.knn_seq(xq, xb, k, metric)
        Db, Ib = flat_ref.knn_blas(xq, xb, k, metric)
        np.testing.assert_allclose(Ds[:, :kk], Db[:, :kk], rtol=1e-4, atol=1e-4)
        d64 = flat_ref.pairwise_f64(xq, xb, metric)
        for r in range(nq):
            for a, b in zip(Is[r, :kk], Ib[r, :kk]):
                assert a == b or abs(d64[r, a] - d64[r, b]) <= 1e-4 * max(abs(d64[r, a]), 1.0)


@settings(max_examples=40, deadline=None)
@given(seed=st.integers(0, 10_000), n=st.integers(0, 40), d=st.integers(1, 40), metric=st.sampled_from([L2, IP]))
def test_index_file_round_trip(seed, n, d, metric):
    x = _arr(seed, n, d)
    blob = faiss_io.write_index_bytes(x, metric)
    assert len(blob) == 45 + 4 * n * d and blob[:4] == (b"IxF2" if metric == L2 else b"IxFI")
    d2, n2, m2, x2 = faiss_io.read_index_bytes(blob)
    assert (d2, n2, m2) == (d, n, metric) and (x2 == x).all()


@settings(max_examples=60, deadline=None)
@given(bits=st.lists(st.booleans(), min_size=0, max_size=70), lo=st.integers(-5, 80), hi=st.integers(-5, 80),
       ids=st.lists(st.integers(-3, 90), max_size=12))
def test_selector_algebra(bits, lo, hi, ids):
    import docqa_b200 as bfk

    n = len(bits)
    m = np.array(bits, dtype=bool)
    to_mask = lambda sel: flat_ref.bitmap_to_mask(sel.to_bitmap(n), n)  # noqa: E731
    a = bfk.IDSelectorBitmap.from_mask(m)
    r = bfk.IDSelectorRange(lo, hi)
    b = bfk.IDSelectorBatch(ids)
    assert (to_mask(a) == m).all()
    want_r = np.array([lo <= i < hi for i in range(n)], dtype=bool)
    want_b = np.array([i in set(ids) for i in range(n)], dtype=bool)
    assert (to_mask(r) == want_r).all() and (to_mask(b) == want_b).all()
    assert (to_mask(bfk.IDSelectorNot(a)) == ~m).all()
    assert (to_mask(bfk.IDSelectorAnd(a, r)) == (m & want_r)).all()
    assert (to_mask(bfk.IDSelectorOr(b, r)) == (want_b | want_r)).all()
    assert to_mask(bfk.IDSelectorAll()).all()
    assert a.to_bitmap(n).size == (n + 7) // 8


@settings(max_examples=30, deadline=None)
@given(seed=st.integers(0, 10_000), n=st.integers(1, 50), d=st.integers(1, 16), k=st.integers(1, 8),
       frac=st.floats(0.0, 1.0), metric=st.sampled_from([L2, IP]))
def test_selected_search_equals_search_over_the_admitted_rows(seed, n, d, k, frac, metric):
    rng = np.random.default_rng(seed)
    xb, xq = _arr(seed, n, d), _arr(seed + 1, 3, d)
    mask = rng.random(n) < frac
    D, I = flat_ref.knn_selected(xq, xb, k, metric, mask)
    ids = np.flatnonzero(mask)
    assert np.isin(I[I >= 0], ids).all() and ((I >= 0).sum(1) == min(k, ids.size)).all()
    if ids.size:
        Dr, Ir = flat_ref.knn(xq, xb[ids], k, metric)
        assert (np.where(Ir >= 0, ids[np.clip(Ir, 0, ids.size - 1)], -1) == I).all()
        np.testing.assert_array_equal(D, Dr)

"""Oracle self-consistency: numpy restatement == C restatement == float64 truth (not gpu)."""
import os

import numpy as np
import pytest

from oracle import cpu_baseline, flat_ref
from tests.parity import check_knn

L2, IP = flat_ref.METRIC_L2, flat_ref.METRIC_INNER_PRODUCT


def unit(rng, n, d):
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x


def test_golden_search_on_reference_fixture(golden_dir, fixture_corpus):
    g = np.load(os.path.join(golden_dir, "docqa_search_golden.npz"))
    q = np.concatenate([fixture_corpus, g["perturbed_queries"]], axis=0)
    assert q.shape == (713, 384)
    # single-query calls with k=3 (llm-qa/main.py:101) and k=5 (BASELINE config 1)
    for r in list(range(0, 649, 37)) + [2, 12, 74, 75, 78, 647, 648] + list(range(649, 713, 9)):
        for k in (3, 5):
            D, I = flat_ref.knn(q[r : r + 1], fixture_corpus, k, L2)
            check_knn(D, I, q[r : r + 1], fixture_corpus, k, L2, I_ref=g["I"][r : r + 1, :k],
                      D_ref64=g["D64"][r : r + 1, :k], tight=True)
            D2, I2 = cpu_baseline.knn(q[r : r + 1], fixture_corpus, k, L2)
            check_knn(D2, I2, q[r : r + 1], fixture_corpus, k, L2, I_ref=g["I"][r : r + 1, :k],
                      D_ref64=g["D64"][r : r + 1, :k], tight=True)
    # a corpus row queried against the corpus finds itself (or its duplicate) at distance 0
    D, I = flat_ref.knn(fixture_corpus[:16], fixture_corpus, 3, L2)
    assert (D[:, 0] == 0).all()
    assert (fixture_corpus[I[:, 0]] == fixture_corpus[:16]).all()
    # batch >= 20 takes the BLAS form: same ids, distances within tolerance of the golden
    D, I = flat_ref.knn(q[:64], fixture_corpus, 5, L2)
    check_knn(D, I, q[:64], fixture_corpus, 5, L2, I_ref=g["I"][:64], D_ref64=g["D64"][:64])
    D, I = cpu_baseline.knn(q[:64], fixture_corpus, 5, L2)
    check_knn(D, I, q[:64], fixture_corpus, 5, L2, I_ref=g["I"][:64], D_ref64=g["D64"][:64])


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("nq", [1, 7, 19, 20, 65])
def test_numpy_and_c_oracles_agree(metric, nq):
    rng = np.random.default_rng(nq * 10 + metric)
    xb, xq = unit(rng, 3000, 96), unit(rng, nq, 96)
    for k in (1, 10, 100):
        Dn, In = flat_ref.knn(xq, xb, k, metric)
        Dc, Ic = cpu_baseline.knn(xq, xb, k, metric)
        check_knn(Dn, In, xq, xb, k, metric)
        check_knn(Dc, Ic, xq, xb, k, metric)
        assert (In == Ic).mean() > 0.999
        np.testing.assert_allclose(Dn, Dc, rtol=1e-5, atol=1e-6)


def test_padding_when_k_exceeds_ntotal():
    rng = np.random.default_rng(1)
    xb, xq = unit(rng, 4, 8), unit(rng, 3, 8)
    for metric, pad in ((L2, np.float32(3.4028235e38)), (IP, np.float32(-3.4028235e38))):
        for fn in (flat_ref.knn, cpu_baseline.knn):
            D, I = fn(xq, xb, 10, metric)
            assert D.shape == (3, 10) and I.dtype == np.int64 and D.dtype == np.float32
            assert (I[:, 4:] == -1).all() and (D[:, 4:] == pad).all()
            assert sorted(I[0, :4].tolist()) == [0, 1, 2, 3]
        D, I = flat_ref.knn(xq, np.zeros((0, 8), np.float32), 3, metric)
        assert (I == -1).all() and (D == pad).all()


def test_ties_resolve_by_id_and_strict_compare():
    d = 16
    rng = np.random.default_rng(2)
    base = unit(rng, 5, d)
    xb = np.concatenate([base, base, base], axis=0)  # every row three times
    for fn in (flat_ref.knn_seq, cpu_baseline.knn_seq):
        D, I = fn(base[:1], xb, 4, L2)
        assert I[0, :3].tolist() == [0, 5, 10] and (D[0, :3] == 0).all()
    D, I = flat_ref.knn_blas(np.repeat(base[:1], 20, 0), xb, 3, L2)
    assert set(I[0].tolist()) == {0, 5, 10}
    assert (D >= 0).all()  # the BLAS form clamps negatives


def test_l2_is_squared_and_ip_is_descending():
    xb = np.array([[0, 0], [3, 4], [1, 0]], np.float32)
    xq = np.array([[0, 0]], np.float32)
    D, I = flat_ref.knn(xq, xb, 3, L2)
    assert I[0].tolist() == [0, 2, 1] and D[0].tolist() == [0.0, 1.0, 25.0]
    D, I = flat_ref.knn(np.array([[1, 1]], np.float32), xb, 3, IP)
    assert I[0].tolist() == [1, 2, 0] and D[0].tolist() == [7.0, 1.0, 0.0]


def test_index_ref_surface():
    rng = np.random.default_rng(3)
    idx = flat_ref.IndexFlatRef(8)
    assert idx.ntotal == 0 and idx.is_trained
    x = unit(rng, 10, 8)
    for i in range(10):  # the reference adds one row per call (indexer.py:41)
        idx.add(x[i : i + 1].astype("float64"))  # .astype('float32') happens inside, like the caller does
    assert idx.ntotal == 10
    with pytest.raises(AssertionError):
        idx.add(np.zeros((1, 7), np.float32))
    with pytest.raises(AssertionError):
        idx.search(x[:1], 0)
    assert (idx.reconstruct(3) == x[3]).all()
    assert idx.remove_ids(np.array([2, 2, 5, 99])) == 2
    assert idx.ntotal == 8 and (idx.reconstruct(2) == x[3]).all()  # later rows shift down
    D, I = idx.search(x[3:4], 1)
    assert I[0, 0] == 2
    other = flat_ref.IndexFlatRef(8)
    other.add(x[:2])
    idx.merge_from(other)
    assert idx.ntotal == 10 and other.ntotal == 0


def test_normalize_l2():
    rng = np.random.default_rng(4)
    x = rng.standard_normal((5, 12)).astype(np.float32)
    x[2] = 0
    y = x.copy()
    flat_ref.normalize_L2(y)
    nr = np.linalg.norm(y, axis=1)
    assert np.allclose(nr[[0, 1, 3, 4]], 1, atol=1e-6) and (y[2] == 0).all()

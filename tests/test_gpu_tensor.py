"""Tensor-core search path (bf16 tcgen05 scan + exact fp32 rerank + certificate). GPU only."""
import numpy as np
import pytest

from oracle import cpu_baseline, flat_ref
from tests.parity import check_knn

pytestmark = pytest.mark.gpu

L2, IP = flat_ref.METRIC_L2, flat_ref.METRIC_INNER_PRODUCT


def unit(rng, n, d):
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x


@pytest.fixture(scope="module")
def bfk():
    import docqa_b200

    assert docqa_b200.get_num_gpus() >= 1
    return docqa_b200


def run(bfk, idx, q, k):
    st = bfk.SearchStats()
    D, I = idx.search(q, k, algo="tensor", stats=st)
    assert st.algo_used == bfk.ALGO_TENSOR
    return D, I, st


@pytest.mark.parametrize("metric", [L2, IP])
def test_tensor_path_on_reference_fixture(bfk, golden_dir, fixture_corpus, metric):
    import os

    g = np.load(os.path.join(golden_dir, "docqa_search_golden.npz"))
    q = np.concatenate([fixture_corpus[:200], g["perturbed_queries"]], axis=0)
    idx = bfk.IndexFlat(384, metric)
    idx.add(fixture_corpus)
    for k in (3, 5, 10):
        D, I, st = run(bfk, idx, q, k)
        check_knn(D, I, q, fixture_corpus, k, metric, tight=True)
    if metric == L2:
        D, I, st = run(bfk, idx, fixture_corpus, 5)  # every row as a query: 23 duplicate groups
        check_knn(D, I, fixture_corpus, fixture_corpus, 5, L2, I_ref=g["I"][:649], D_ref64=g["D64"][:649], tight=True)
        assert (D[:, 0] == 0).all()


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d", [384, 100, 64, 37])
@pytest.mark.parametrize("nq", [9, 64, 128, 129, 300])
def test_tensor_path_random(bfk, metric, d, nq):
    rng = np.random.default_rng(1000 * d + nq + metric)
    n = 20011
    xb, xq = unit(rng, n, d), unit(rng, nq, d)
    idx = bfk.IndexFlat(d, metric)
    idx.add(xb)
    for k in (1, 10, 24):
        D, I, st = run(bfk, idx, xq, k)
        check_knn(D, I, xq, xb, k, metric, tight=True)
        De, Ie = idx.search(xq, k, algo="exact")
        assert (I == Ie).all()
        np.testing.assert_allclose(D, De, rtol=1e-5, atol=1e-6)


def test_tensor_path_small_and_ragged_corpora(bfk):
    rng = np.random.default_rng(3)
    for n in (1, 5, 255, 256, 257, 1000):
        xb, xq = unit(rng, n, 384), unit(rng, 17, 384)
        idx = bfk.IndexFlatL2(384)
        idx.add(xb)
        D, I, st = run(bfk, idx, xq, 10)
        check_knn(D, I, xq, xb, 10, L2, tight=True)


def test_tensor_path_unnormalised_rows(bfk):
    rng = np.random.default_rng(4)
    xb = (rng.standard_normal((30000, 128)) * rng.uniform(0.1, 5.0, size=(30000, 1))).astype(np.float32)
    xq = (rng.standard_normal((40, 128)) * 2.0).astype(np.float32)
    for metric in (L2, IP):
        idx = bfk.IndexFlat(128, metric)
        idx.add(xb)
        D, I, st = run(bfk, idx, xq, 10)
        check_knn(D, I, xq, xb, 10, metric)


def test_certificate_fallback_on_ties_and_adversarial_order(bfk):
    rng = np.random.default_rng(5)
    d = 384
    # (a) all rows identical: nothing can be certified -> exact re-run on the GPU, smallest ids win
    same = np.repeat(unit(rng, 1, d), 5000, axis=0)
    idx = bfk.IndexFlatL2(d)
    idx.add(same)
    q = unit(rng, 12, d)
    D, I, st = run(bfk, idx, q, 10)
    assert st.n_fallback == 12
    assert (I == np.arange(10)[None, :]).all() and np.ptp(D, axis=1).max() == 0
    # (b) 1% duplicates + farthest-first order for query 0
    xb = unit(rng, 40000, d)
    src = rng.integers(0, 40000, size=400)
    xb[rng.integers(0, 40000, size=400)] = xb[src]
    q = unit(rng, 33, d)
    xb = np.ascontiguousarray(xb[np.argsort(-((xb - q[0]) ** 2).sum(1))])
    idx2 = bfk.IndexFlatL2(d)
    idx2.add(xb)
    D, I, st = run(bfk, idx2, q, 10)
    check_knn(D, I, q, xb, 10, L2, tight=True)
    # (c) queries that are corpus rows: distance exactly 0 first
    D, I, st = run(bfk, idx2, xb[100:120], 5)
    assert (D[:, 0] == 0).all()
    check_knn(D, I, xb[100:120], xb, 5, L2, tight=True)


def test_tensor_path_bf16_storage(bfk):
    import torch

    rng = np.random.default_rng(6)
    xb, xq = unit(rng, 50000, 384), unit(rng, 70, 384)
    xb16 = torch.from_numpy(xb).to(torch.bfloat16).to(torch.float32).numpy()
    for metric in (L2, IP):
        idx = bfk.IndexFlat(384, metric, storage="bf16")
        idx.add(xb)
        D, I, st = run(bfk, idx, xq, 10)
        check_knn(D, I, xq, xb16, 10, metric, tight=True)


def test_tensor_path_limits_and_auto_dispatch(bfk):
    rng = np.random.default_rng(7)
    xb, xq = unit(rng, 3000, 384), unit(rng, 40, 384)
    idx = bfk.IndexFlatL2(384)
    idx.add(xb)
    with pytest.raises(RuntimeError):
        idx.search(xq, 200, algo="tensor")  # k > 128 is served by the exact path only
    st = bfk.SearchStats()
    D, I = idx.search(xq, 200, stats=st)
    assert st.algo_used == bfk.ALGO_EXACT_SIMT
    check_knn(D, I, xq, xb, 200, L2, tight=True)
    D, I = idx.search(xq, 10, stats=st)
    assert st.algo_used == bfk.ALGO_TENSOR
    # a handful of queries over a small fp32 corpus: the fp32 scan's lower fixed cost wins
    D, I = idx.search(xq[:8], 10, stats=st)
    assert st.algo_used == bfk.ALGO_EXACT_SIMT
    check_knn(D, I, xq[:8], xb, 10, L2, tight=True)
    # over a large one the bf16 scan copy is half the bytes and the answers are certified: tensor path, even for nq = 1
    big = unit(rng, 180_000, 384)
    idx2 = bfk.IndexFlatL2(384)
    idx2.add(big)
    for nq in (1, 8):
        D, I = idx2.search(xq[:nq], 10, stats=st)
        assert st.algo_used == bfk.ALGO_TENSOR
        check_knn(D, I, xq[:nq], big, 10, L2, tight=True)
    # bf16-stored rows have no fp32 copy to scan faster: tensor path at any size
    idx3 = bfk.IndexFlat(384, L2, storage="bf16")
    idx3.add(xb)
    D, I = idx3.search(xq[:2], 10, stats=st)
    assert st.algo_used == bfk.ALGO_TENSOR
    D, I = idx.search(xq[:8], 10, stats=st, algo="exact")
    assert st.algo_used == bfk.ALGO_EXACT_SIMT


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("storage", ["f32", "bf16"])
def test_tensor_path_k100_and_d768(bfk, metric, storage):
    """BASELINE config 3 in miniature: 768-d rows (query tile streamed, not resident), k = 100."""
    import torch

    rng = np.random.default_rng(8)
    xb, xq = unit(rng, 60000, 768), unit(rng, 300, 768)
    ref = xb if storage == "f32" else torch.from_numpy(xb).to(torch.bfloat16).to(torch.float32).numpy()
    idx = bfk.IndexFlat(768, metric, storage=storage)
    idx.add(xb)
    for k in (10, 100, 128):
        D, I, st = run(bfk, idx, xq, k)
        check_knn(D, I, xq, ref, k, metric, tight=True)
        assert st.n_fallback <= 3, (k, st.n_fallback)
    # 384-d, k = 100 on a batch of one query tile
    xb2, xq2 = unit(rng, 40000, 384), unit(rng, 50, 384)
    idx2 = bfk.IndexFlat(384, metric)
    idx2.add(xb2)
    D, I, st = run(bfk, idx2, xq2, 100)
    check_knn(D, I, xq2, xb2, 100, metric, tight=True)


def test_tensor_path_full_size_1m_x_384(bfk):
    """BASELINE config 2 at full size, batches 64 and 1024."""
    import torch

    n, d, k = 1_000_000, 384, 10
    g = torch.Generator(device="cuda").manual_seed(1234)
    xb = torch.randn((n, d), generator=g, device="cuda", dtype=torch.float32)
    xb /= xb.norm(dim=1, keepdim=True)
    g2 = torch.Generator(device="cuda").manual_seed(4321)
    q = torch.randn((1024, d), generator=g2, device="cuda", dtype=torch.float32)
    q /= q.norm(dim=1, keepdim=True)
    idx = bfk.IndexFlatL2(d)
    idx.add(xb)
    st = bfk.SearchStats()
    D, I = idx.search(q, k, algo="tensor", stats=st)
    assert st.n_fallback == 0  # random unit vectors certify
    D, I = D.cpu().numpy(), I.cpu().numpy()
    xb_h, q_h = xb.cpu().numpy(), q.cpu().numpy()
    Dc, Ic = cpu_baseline.knn(q_h[:64], xb_h, k, L2)  # BLAS form of the CPU restatement
    check_knn(D[:64], I[:64], q_h[:64], xb_h, k, L2, I_ref=Ic, D_ref64=Dc.astype(np.float64))
    # size-independent properties on all 1024 rows
    assert (np.diff(D, axis=1) >= 0).all() and I.min() >= 0 and I.max() < n
    rec = ((xb_h[I.reshape(-1)].reshape(1024, k, d) - q_h[:, None, :]) ** 2).sum(-1)
    np.testing.assert_allclose(D, rec, rtol=1e-3, atol=1e-6)
    # a 64-query call and the exact path agree with the 1024-query call
    D64, I64 = idx.search(q[:64], k, algo="tensor")
    assert (I64.cpu().numpy() == I[:64]).all()
    De, Ie = idx.search(q[:16], k, algo="exact")
    assert (Ie.cpu().numpy() == I[:16]).all()


def test_config3_full_size_10m_x_768_bf16_ip(bfk):
    """BASELINE config 3 at full size: bf16-stored 10M x 768, IndexFlatIP, batch 4096, k = 100.

    The CPU oracle cannot hold this corpus; the tensor path is checked against the independent
    exact fp32-accumulate scan (K3a) on a query subsample, plus size-independent properties.
    """
    import torch

    n, d, k, nq = 10_000_000, 768, 100, 4096
    idx = bfk.IndexFlatIP(d, storage="bf16")
    idx.reserve(n)
    g = torch.Generator(device="cuda").manual_seed(1234)
    for r0 in range(0, n, 1 << 20):
        m = min(1 << 20, n - r0)
        x = torch.randn((m, d), generator=g, device="cuda", dtype=torch.float32)
        x /= x.norm(dim=1, keepdim=True)
        idx.add(x)
        del x
    assert idx.ntotal == n
    g2 = torch.Generator(device="cuda").manual_seed(4321)
    q = torch.randn((nq, d), generator=g2, device="cuda", dtype=torch.float32)
    q /= q.norm(dim=1, keepdim=True)
    st = bfk.SearchStats()
    D, I = idx.search(q, k, algo="tensor", stats=st)
    assert st.algo_used == bfk.ALGO_TENSOR and st.n_fallback <= nq // 100, st.n_fallback
    De, Ie = idx.search(q[:16], k, algo="exact")
    assert torch.equal(I[:16], Ie)
    torch.testing.assert_close(D[:16], De, rtol=1e-5, atol=1e-6)
    Dh, Ih = D.cpu().numpy(), I.cpu().numpy()
    assert (np.diff(Dh, axis=1) <= 0).all()  # inner products, best (largest) first
    assert Ih.min() >= 0 and Ih.max() < n
    assert all(len(set(r)) == k for r in Ih[::64].tolist())
    # distances recomputable from the stored (bf16) rows
    rows = torch.from_numpy(idx.reconstruct_n(int(Ih[5, 0]), 1)).cuda()
    assert abs(float((rows[0] * q[5]).sum()) - float(Dh[5, 0])) < 1e-5


# ---- the certificate's error bound at large d ("Exactness") --------------------------
# The margin K5 re-scores inside grows with the row length (tensor-pipe accumulation over ld16 terms); these
# cases sit at the d where the old constant was not proved: 1024 < d <= BFK_MAX_D.

@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("storage", ["f32", "bf16"])
@pytest.mark.parametrize("d", [1024, 1536, 2048])
def test_certificate_large_d_random_rows(bfk, d, storage, metric):
    import torch
    from tests.conftest import note

    rng = np.random.default_rng(d + metric)
    xb, xq = unit(rng, 20011, d), unit(rng, 70, d)
    stored = torch.from_numpy(xb).to(torch.bfloat16).to(torch.float32).numpy() if storage == "bf16" else xb
    idx = bfk.IndexFlat(d, metric, storage=storage)
    idx.add(xb)
    D, I, st = run(bfk, idx, xq, 10)
    note(f"large_d_random d={d} storage={storage} metric={metric} n_fallback={st.n_fallback} second_pass={st.n_candidates}")
    check_knn(D, I, xq, stored, 10, metric, tight=True)
    De, Ie = idx.search(xq, 10, algo="exact")
    assert (I == Ie).all()
    np.testing.assert_allclose(D, De, rtol=1e-5, atol=1e-6)
    assert st.n_fallback <= 8  # i.i.d. rows are certifiable at every d: the bound must not swallow the answer


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d", [1024, 2048])
def test_certificate_large_d_all_positive_rows_and_wide_norms(bfk, d, metric):
    """Worst case for the accumulation bound: every product q_i * x_i is positive, so sum|q_i x_i| = <q, x>,
    and the row norms span three decades (the bound is relative to ||q|| max||x||)."""
    from tests.conftest import note

    rng = np.random.default_rng(7 * d + metric)
    n, nq, k = 20011, 40, 10
    xb = np.abs(rng.standard_normal((n, d))).astype(np.float32)
    xb *= (10.0 ** rng.uniform(-1.5, 1.5, size=(n, 1)) / np.linalg.norm(xb, axis=1, keepdims=True)).astype(np.float32)
    xq = np.abs(rng.standard_normal((nq, d))).astype(np.float32)
    xq /= np.linalg.norm(xq, axis=1, keepdims=True)
    idx = bfk.IndexFlat(d, metric)
    idx.add(xb)
    D, I, st = run(bfk, idx, xq, k)
    note(f"large_d_positive d={d} metric={metric} n_fallback={st.n_fallback} second_pass={st.n_candidates}")
    check_knn(D, I, xq, xb, k, metric)
    De, Ie = idx.search(xq, k, algo="exact")
    assert (I == Ie).all()
    np.testing.assert_allclose(D, De, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("d", [1024, 1536, 2048])
def test_certificate_large_d_near_ties_fall_back_to_the_exact_scan(bfk, d):
    """A cluster of 700 rows whose exact distances to the query differ by far less than the margin: more rows
    sit inside the cut than K5 re-scores (512), so nothing can be certified there; the answer must then come
    from the fp32 scan on the GPU (n_fallback > 0) and equal it."""
    from tests.conftest import note

    rng = np.random.default_rng(11 * d)
    n, k = 20000, 10
    xb = unit(rng, n, d)
    v = unit(rng, 1, d)
    where = rng.choice(n, size=700, replace=False)
    xb[where] = v + 1e-3 * unit(rng, 700, d)
    xq = np.concatenate([v + 0.05 * unit(rng, 6, d), unit(rng, 20, d)], axis=0).astype(np.float32)
    idx = bfk.IndexFlatL2(d)
    idx.add(xb)
    D, I, st = run(bfk, idx, xq, k)
    note(f"large_d_near_ties d={d} n_fallback={st.n_fallback} second_pass={st.n_candidates}")
    assert st.n_fallback >= 6  # the six queries next to the cluster
    check_knn(D, I, xq, xb, k, L2)
    De, Ie = idx.search(xq, k, algo="exact")
    assert (I == Ie).all()  # the fall-back IS the exact scan, so even inside the near-ties the ids agree
    np.testing.assert_allclose(D, De, rtol=1e-5, atol=1e-6)
    assert set(I[:6].reshape(-1).tolist()) <= set(where.tolist())

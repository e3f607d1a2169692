"""Selector-restricted search (faiss SearchParameters(sel=...)) and range_search against the CPU oracle. GPU only."""
import numpy as np
import pytest

from oracle import flat_ref
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

    assert docqa_b200.get_num_gpus() >= 1, "GPU tests need a CUDA device"
    return docqa_b200


def check_selected(D, I, xq, xb, k, metric, mask):
    """(D, I) must be the exact top-k of the admitted rows, with original row ids."""
    ids = np.flatnonzero(mask)
    assert np.isin(I[I >= 0], ids).all(), "a rejected row was returned"
    # map ids back to positions in the admitted subset and reuse the standard comparator
    pos = np.full(xb.shape[0] + 1, -1, dtype=np.int64)
    pos[ids] = np.arange(ids.size)
    I_sub = np.where(I >= 0, pos[np.clip(I, 0, xb.shape[0])], -1)
    check_knn(D, I_sub, xq, xb[ids], k, metric)


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("algo", ["exact", "tensor"])
@pytest.mark.parametrize("frac", [0.5, 0.02])
def test_search_with_bitmap_selector(bfk, metric, algo, frac):
    rng = np.random.default_rng(7)
    n, d, k = 20_000, 96, 10
    xb, xq = unit(rng, n, d), unit(rng, 33 if algo == "tensor" else 5, d)
    mask = rng.random(n) < frac
    idx = bfk.IndexFlat(d, metric)
    idx.add(xb)
    sel = bfk.IDSelectorBitmap.from_mask(mask)
    D, I = idx.search(xq, k, params=bfk.SearchParameters(sel=sel), algo=algo)
    check_selected(D, I, xq, xb, k, metric, mask)
    # the oracle mirror agrees position by position outside ties
    ref = flat_ref.IndexFlatRef(d, metric)
    ref.add(xb)
    Dr, Ir = ref.search(xq, k, params=bfk.SearchParameters(sel=sel))
    assert (I == Ir).mean() > 0.99
    # and an unrestricted search afterwards is unaffected by the selector state
    D2, I2 = idx.search(xq, k, algo=algo)
    check_knn(D2, I2, xq, xb, k, metric)


def test_selector_kinds_and_padding(bfk):
    rng = np.random.default_rng(8)
    n, d = 5_000, 64
    xb, xq = unit(rng, n, d), unit(rng, 4, d)
    idx = bfk.IndexFlatL2(d)
    idx.add(xb)
    ref = flat_ref.IndexFlatRef(d, L2)
    ref.add(xb)
    sels = [
        bfk.IDSelectorRange(100, 900),
        bfk.IDSelectorBatch([5, 17, 4999, 123456, -3]),          # out-of-range ids are ignored
        bfk.IDSelectorNot(bfk.IDSelectorRange(0, 4990)),
        bfk.IDSelectorAnd(bfk.IDSelectorRange(0, 2000), bfk.IDSelectorNot(bfk.IDSelectorRange(10, 1990))),
        bfk.IDSelectorOr(bfk.IDSelectorArray([1, 2, 3]), bfk.IDSelectorRange(4000, 4003)),
        bfk.IDSelectorAll(),
    ]
    for sel in sels:
        p = bfk.SearchParameters(sel=sel)
        mask = flat_ref.bitmap_to_mask(sel.to_bitmap(n), n)
        for algo in ("exact", "tensor"):
            D, I = idx.search(xq, 8, params=p, algo=algo)
            Dr, Ir = ref.search(xq, 8, params=p)
            check_selected(D, I, xq, xb, 8, L2, mask)
            assert ((I == -1) == (Ir == -1)).all()          # fewer than k admitted rows -> -1 padding
    # nothing admitted at all
    D, I = idx.search(xq, 3, params=bfk.SearchParameters(sel=bfk.IDSelectorRange(7, 7)))
    assert (I == -1).all() and (D == np.finfo(np.float32).max).all()


def test_selector_with_torch_tensors_and_bf16_storage(bfk):
    import torch

    rng = np.random.default_rng(9)
    n, d, k = 30_000, 128, 5
    xb, xq = unit(rng, n, d), unit(rng, 16, d)
    mask = rng.random(n) < 0.1
    idx = bfk.IndexFlat(d, IP, storage="bf16")
    idx.add(xb)
    xs = idx.reconstruct_n(0, n)  # the bf16-rounded rows the index really holds
    q = torch.from_numpy(xq).cuda()
    D, I = idx.search(q, k, params=bfk.SearchParameters(sel=bfk.IDSelectorBitmap.from_mask(mask)))
    assert D.is_cuda and I.is_cuda
    check_selected(D.cpu().numpy(), I.cpu().numpy(), xq, xs, k, IP, mask)


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("storage", ["f32", "bf16"])
def test_range_search_matches_oracle(bfk, metric, storage):
    rng = np.random.default_rng(10)
    n, d = 12_000, 48
    xb, xq = unit(rng, n, d), unit(rng, 11, d)
    idx = bfk.IndexFlat(d, metric, storage=storage)
    idx.add(xb)
    xs = idx.reconstruct_n(0, n)
    radius = 1.55 if metric == L2 else 0.22
    lims, D, I = idx.range_search(xq, radius)
    lr, Dr, Ir = flat_ref.range_search(xq, xs, radius, metric)
    assert lims.dtype == np.int64 and D.dtype == np.float32 and I.dtype == np.int64
    assert lims[0] == 0 and lims[-1] == D.size == I.size and lims[-1] > 50
    d64 = flat_ref.pairwise_f64(xq, xs, metric)
    for i in range(xq.shape[0]):
        got, want = I[lims[i] : lims[i + 1]], Ir[lr[i] : lr[i + 1]]
        # rows may differ only where the distance sits within rounding of the radius
        diff = np.setxor1d(got, want)
        assert (np.abs(d64[i, diff] - radius) < 1e-5).all()
        assert (np.diff(got) > 0).all()  # id order, no duplicates
        np.testing.assert_allclose(D[lims[i] : lims[i + 1]], d64[i, got], rtol=1e-4, atol=1e-6)
    # empty result and empty index
    lims, D, I = idx.range_search(xq, -1.0 if metric == L2 else 5.0)
    assert (lims == 0).all() and D.size == 0 and I.size == 0
    empty = bfk.IndexFlat(d, metric)
    lims, D, I = empty.range_search(xq, radius)
    assert (lims == 0).all() and D.size == 0

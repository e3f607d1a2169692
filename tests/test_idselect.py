"""Selector classes (faiss IDSelector* mirror) and the oracle's selector / range search: CPU only."""
import numpy as np

from oracle import flat_ref


def _pkg():
    import docqa_b200

    return docqa_b200


def test_bitmap_layout_matches_faiss():
    bfk = _pkg()
    # faiss IDSelectorBitmap: id i is a member iff bitmap[i >> 3] & (1 << (i & 7))
    sel = bfk.IDSelectorBatch([0, 9, 23])
    bm = sel.to_bitmap(24)
    assert bm.dtype == np.uint8 and bm.tolist() == [0b00000001, 0b00000010, 0b10000000]
    assert sel.is_member(9) and not sel.is_member(8)
    assert bfk.IDSelectorRange(3, 6).to_bitmap(10).tolist() == [0b00111000, 0]
    assert bfk.IDSelectorNot(bfk.IDSelectorRange(0, 8)).to_bitmap(10)[0] == 0
    m = np.array([1, 0, 1, 1, 0, 0, 0, 0, 1], dtype=bool)
    assert (flat_ref.bitmap_to_mask(bfk.IDSelectorBitmap.from_mask(m).to_bitmap(9), 9) == m).all()
    both = bfk.IDSelectorAnd(bfk.IDSelectorRange(0, 5), bfk.IDSelectorRange(3, 9))
    assert flat_ref.bitmap_to_mask(both.to_bitmap(9), 9).nonzero()[0].tolist() == [3, 4]
    either = bfk.IDSelectorOr(bfk.IDSelectorArray([1]), bfk.IDSelectorArray([7]))
    assert flat_ref.bitmap_to_mask(either.to_bitmap(9), 9).nonzero()[0].tolist() == [1, 7]


def test_oracle_selected_search_and_range_search():
    bfk = _pkg()
    rng = np.random.default_rng(0)
    xb = rng.standard_normal((300, 16)).astype(np.float32)
    xq = rng.standard_normal((4, 16)).astype(np.float32)
    ref = flat_ref.IndexFlatRef(16)
    ref.add(xb)
    sel = bfk.IDSelectorRange(50, 60)
    D, I = ref.search(xq, 12, params=bfk.SearchParameters(sel=sel))
    assert ((I[:, :10] >= 50) & (I[:, :10] < 60)).all() and (I[:, 10:] == -1).all()
    Df, If = flat_ref.knn(xq, xb[50:60], 10, flat_ref.METRIC_L2)
    assert (I[:, :10] == If + 50).all() and np.allclose(D[:, :10], Df)
    lims, Dr, Ir = ref.range_search(xq, 20.0)
    full = flat_ref.pairwise_seq(xq, xb, flat_ref.METRIC_L2)
    for i in range(4):
        assert Ir[lims[i] : lims[i + 1]].tolist() == np.flatnonzero(full[i] < 20.0).tolist()

"""faiss.SearchParameters / IDSelector* for the flat index (LangChain-side extras).

faiss lets `index.search(x, k, params=SearchParameters(sel=IDSelector...))` restrict the rows a search may
return.  The reference does not use it today; the patient-snippets endpoint synthese-comparative expects
(core/retrieval_client.py:58-91: patient_id + date window) is a search over the rows of one patient, which
is exactly this.  Every selector lowers to faiss's IDSelectorBitmap layout -- bit i of byte i >> 3 set <=>
row i admitted -- and that bitmap is what crosses the C ABI (bfk_search_sel).
"""
from __future__ import annotations

import numpy as np


class IDSelector:
    """Base class; subclasses say which row ids a search may return."""

    def to_bitmap(self, ntotal: int) -> np.ndarray:
        raise NotImplementedError

    def is_member(self, i: int) -> bool:
        bm = self.to_bitmap(int(i) + 1)
        return bool((bm[int(i) >> 3] >> (int(i) & 7)) & 1)


def _bitmap_from_bool(mask: np.ndarray) -> np.ndarray:
    return np.packbits(np.ascontiguousarray(mask, dtype=bool), bitorder="little")


class IDSelectorAll(IDSelector):
    def to_bitmap(self, ntotal):
        return _bitmap_from_bool(np.ones(ntotal, dtype=bool))


class IDSelectorRange(IDSelector):
    """ids in [imin, imax) -- faiss.IDSelectorRange(imin, imax)."""

    def __init__(self, imin: int, imax: int, assume_sorted: bool = False):
        self.imin, self.imax = int(imin), int(imax)

    def to_bitmap(self, ntotal):
        m = np.zeros(ntotal, dtype=bool)
        m[max(self.imin, 0):max(min(self.imax, ntotal), 0)] = True
        return _bitmap_from_bool(m)


class IDSelectorArray(IDSelector):
    """an explicit id list -- faiss.IDSelectorArray(ids) / IDSelectorBatch(ids)."""

    def __init__(self, ids):
        self.ids = np.asarray(ids, dtype=np.int64).reshape(-1)

    def to_bitmap(self, ntotal):
        m = np.zeros(ntotal, dtype=bool)
        ids = self.ids[(self.ids >= 0) & (self.ids < ntotal)]
        m[ids] = True
        return _bitmap_from_bool(m)


class IDSelectorBatch(IDSelectorArray):
    pass


class IDSelectorBitmap(IDSelector):
    """faiss.IDSelectorBitmap(bitmap): uint8 array, bit i of byte i >> 3; ids past the array are rejected.

    Also accepts a boolean mask over rows (`IDSelectorBitmap.from_mask`).
    """

    def __init__(self, bitmap):
        self.bitmap = np.ascontiguousarray(bitmap, dtype=np.uint8).reshape(-1)

    @classmethod
    def from_mask(cls, mask):
        return cls(_bitmap_from_bool(np.asarray(mask, dtype=bool).reshape(-1)))

    def to_bitmap(self, ntotal):
        nb = (ntotal + 7) // 8
        out = np.zeros(nb, dtype=np.uint8)
        m = min(nb, self.bitmap.size)
        out[:m] = self.bitmap[:m]
        return out


class IDSelectorNot(IDSelector):
    def __init__(self, sel: IDSelector):
        self.sel = sel

    def to_bitmap(self, ntotal):
        return np.bitwise_not(self.sel.to_bitmap(ntotal))


class IDSelectorAnd(IDSelector):
    def __init__(self, lhs: IDSelector, rhs: IDSelector):
        self.lhs, self.rhs = lhs, rhs

    def to_bitmap(self, ntotal):
        return np.bitwise_and(self.lhs.to_bitmap(ntotal), self.rhs.to_bitmap(ntotal))


class IDSelectorOr(IDSelector):
    def __init__(self, lhs: IDSelector, rhs: IDSelector):
        self.lhs, self.rhs = lhs, rhs

    def to_bitmap(self, ntotal):
        return np.bitwise_or(self.lhs.to_bitmap(ntotal), self.rhs.to_bitmap(ntotal))


class SearchParameters:
    """faiss.SearchParameters(sel=None)."""

    def __init__(self, sel: IDSelector | None = None):
        self.sel = sel

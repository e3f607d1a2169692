"""faiss.IndexFlat* drop-in for the DocQA semantic-indexer, backed by libbfk.so on a B200.

Mirrors the slice of the faiss Python surface the reference touches:

    semantic-indexer/indexer.py:39,104   faiss.IndexFlatL2(dimension)
    semantic-indexer/indexer.py:41       index.add(np.float32[n, d])
    semantic-indexer/indexer.py:109      index.ntotal
    semantic-indexer/indexer.py:27       faiss.write_index(index, path)
    semantic-indexer/indexer.py:99       faiss.read_index(path)
    llm-qa/main.py:35,59                 faiss.read_index / raw_index.ntotal
    llm-qa/main.py:53-58,101,117         langchain_community FAISS(index=...) ->
                                         index.search(np.float32[1, d], k) plus the side
                                         paths reconstruct / remove_ids / merge_from

Names, argument meaning and error behaviour follow faiss: shape / k checks are Python
`assert`s (AssertionError), library failures are RuntimeError, (D, I) are new
float32 / int64 C-contiguous arrays of shape [n, k], missing results are
I = -1, D = +/-3.4028235e38.  Extension: CUDA torch tensors are accepted for `add` and
`search` and then stay on the device (zero copy, results are CUDA tensors).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from . import _lib
from ._lib import (ALGO_AUTO, ALGO_EXACT_SIMT, ALGO_TENSOR, METRIC_INNER_PRODUCT, METRIC_L2,
                   STORE_BF16, STORE_F32, SearchStats)

_STORAGE = {"f32": STORE_F32, "fp32": STORE_F32, "float32": STORE_F32,
            "bf16": STORE_BF16, "bfloat16": STORE_BF16, STORE_F32: STORE_F32, STORE_BF16: STORE_BF16}
_ALGO = {None: ALGO_AUTO, "auto": ALGO_AUTO, "exact": ALGO_EXACT_SIMT, "simt": ALGO_EXACT_SIMT,
         "tensor": ALGO_TENSOR, ALGO_AUTO: ALGO_AUTO, ALGO_EXACT_SIMT: ALGO_EXACT_SIMT,
         ALGO_TENSOR: ALGO_TENSOR}


def _default_device() -> int:
    return int(os.environ.get("BFK_DEVICE", "0"))


def _is_torch(x) -> bool:
    return type(x).__module__.split(".")[0] == "torch"


def _np_ptr(a: np.ndarray) -> ctypes.c_void_p:
    return ctypes.c_void_p(a.ctypes.data)


def _torch_stream(device: int):
    """The current CUDA stream of torch on `device`, as the integer handle the C ABI takes (void*)."""
    import torch

    try:  # the raw handle without building a torch.cuda.Stream object: ~0.2 us instead of ~1.5 us per search
        return torch._C._cuda_getCurrentRawStream(device)
    except AttributeError:
        return torch.cuda.current_stream(device).cuda_stream


class IndexFlat:
    """Exact brute-force index over fp32 vectors (faiss.IndexFlat)."""

    def __init__(self, d: int, metric: int = METRIC_L2, *, device: int | None = None,
                 storage="f32", _handle=None):
        self._L = _lib.load()
        self.verbose = False
        self.is_trained = True
        if _handle is not None:
            self._h = _handle
            return
        h = ctypes.c_void_p()
        dev = _default_device() if device is None else int(device)
        _lib.check(self._L.bfk_create(int(d), int(metric), dev, _STORAGE[storage], ctypes.byref(h)))
        self._h = h

    # -- attributes faiss exposes ------------------------------------------------
    @property
    def d(self) -> int:
        d = self.__dict__.get("_d")  # fixed at construction: one ctypes call per index, not one per search
        if d is None:
            d = self.__dict__["_d"] = int(self._L.bfk_dim(self._h))
        return d

    @property
    def ntotal(self) -> int:
        return int(self._L.bfk_ntotal(self._h))

    @property
    def metric_type(self) -> int:
        return int(self._L.bfk_metric(self._h))

    @property
    def code_size(self) -> int:
        return 4 * self.d

    @property
    def device(self) -> int:
        dev = self.__dict__.get("_device")
        if dev is None:
            dev = self.__dict__["_device"] = int(self._L.bfk_device(self._h))
        return dev

    @property
    def storage(self) -> str:
        return "bf16" if self._L.bfk_storage(self._h) == STORE_BF16 else "f32"

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._L.bfk_free(h)
            except Exception:
                pass

    # -- mutation ---------------------------------------------------------------------
    def reserve(self, n: int) -> None:
        """Pre-size corpus storage for n rows (extension; avoids regrowth copies)."""
        _lib.check(self._L.bfk_reserve(self._h, int(n)))

    def add(self, x) -> None:
        n, d = x.shape
        assert d == self.d
        if _is_torch(x):
            import torch

            assert x.is_cuda and x.device.index == self.device, "torch input must live on the index's GPU"
            x = x.to(torch.float32).contiguous()
            _lib.check(self._L.bfk_add(self._h, ctypes.c_void_p(x.data_ptr()), n, 1, _torch_stream(self.device)))
            return
        x = np.ascontiguousarray(x, dtype="float32")
        _lib.check(self._L.bfk_add(self._h, _np_ptr(x), n, 0, None))

    def add_with_ids(self, x, ids):
        raise RuntimeError("add_with_ids not implemented for this type of index")

    def add_pooled(self, token_embeddings, attention_mask) -> None:
        """Fused Pooling(mean)+Normalize+add of encoder outputs [B, T, d] (CUDA tensors).

        The device-resident replacement for indexer.py:37 (tail of encode) + :41 (add).
        """
        from .pooling import _tok_dtype, _mask_dtype

        tok = token_embeddings.contiguous()
        mask = attention_mask.contiguous()
        B, T, H = tok.shape
        assert H == self.d
        assert tuple(mask.shape) == (B, T)
        assert tok.is_cuda and mask.is_cuda and tok.device.index == self.device
        _lib.check(self._L.bfk_add_pooled(self._h, ctypes.c_void_p(tok.data_ptr()), _tok_dtype(tok),
                                           ctypes.c_void_p(mask.data_ptr()), _mask_dtype(mask), B, T, H,
                                           _torch_stream(self.device)))

    def reset(self) -> None:
        _lib.check(self._L.bfk_reset(self._h))

    def remove_ids(self, ids) -> int:
        ids = np.ascontiguousarray(np.asarray(ids).reshape(-1), dtype="int64")
        nrem = ctypes.c_int64(0)
        _lib.check(self._L.bfk_remove_ids(self._h, _np_ptr(ids), ids.size, ctypes.byref(nrem)))
        return int(nrem.value)

    def merge_from(self, other: "IndexFlat", add_id: int = 0) -> None:
        assert other.d == self.d
        if other.metric_type != self.metric_type:
            raise RuntimeError("merge_from: metric types differ")
        if other.ntotal:
            self.add(other.reconstruct_n(0, other.ntotal))
        other.reset()

    # -- queries ------------------------------------------------------------------------
    def search(self, x, k, *, params=None, D=None, I=None, algo=None, id_offset: int = 0, stats=None,
               defer: bool = False):
        n, d = x.shape
        assert d == self.d
        assert k > 0
        k = int(k)
        st = stats if isinstance(stats, SearchStats) else None
        stp = ctypes.byref(st) if st is not None else None
        sel = getattr(params, "sel", None) if params is not None else None
        if sel is not None:
            return self._search_sel(x, k, sel, D, I, algo, id_offset, stp)
        if _is_torch(x):
            import torch

            assert x.is_cuda and x.device.index == self.device, "torch input must live on the index's GPU"
            xq = x if (x.dtype is torch.float32 and x.is_contiguous()) else x.to(torch.float32).contiguous()
            if D is None:
                D = torch.empty((n, k), dtype=torch.float32, device=x.device)
            else:
                assert D.shape == (n, k) and D.dtype is torch.float32 and D.is_contiguous()
            if I is None:
                I = torch.empty((n, k), dtype=torch.int64, device=x.device)
            else:
                assert I.shape == (n, k) and I.dtype is torch.int64 and I.is_contiguous()
            # plain integers for the pointer arguments: ctypes converts them itself (argtypes are c_void_p)
            _lib.check(self._L.bfk_search_ex(self._h, xq.data_ptr(), n, k, D.data_ptr(), I.data_ptr(), 1,
                                              _torch_stream(self.device),
                                              _ALGO[algo] | (_lib.ALGO_FLAG_DEFER if defer else 0), int(id_offset), stp))
            return D, I
        x = np.ascontiguousarray(x, dtype="float32")
        if D is None:
            D = np.empty((n, k), dtype=np.float32)
        else:
            assert D.shape == (n, k)
            assert D.dtype == np.float32 and D.flags.c_contiguous
        if I is None:
            I = np.empty((n, k), dtype=np.int64)
        else:
            assert I.shape == (n, k)
            assert I.dtype == np.int64 and I.flags.c_contiguous
        # plain integer addresses: ctypes converts them for the c_void_p parameters itself
        _lib.check(self._L.bfk_search_ex(self._h, x.ctypes.data, n, k, D.ctypes.data, I.ctypes.data, 0, None,
                                          _ALGO[algo], int(id_offset), stp))
        return D, I

    def _search_sel(self, x, k, sel, D, I, algo, id_offset, stp):
        """search restricted to the rows `sel` admits (faiss SearchParameters(sel=...)): bfk_search_sel."""
        n = x.shape[0]
        bm = np.ascontiguousarray(sel.to_bitmap(self.ntotal), dtype=np.uint8)
        if bm.size == 0:
            bm = np.zeros(1, dtype=np.uint8)
        if _is_torch(x):
            import torch

            assert x.is_cuda and x.device.index == self.device, "torch input must live on the index's GPU"
            xq = x.to(torch.float32).contiguous()
            D = torch.empty((n, k), dtype=torch.float32, device=x.device) if D is None else D
            I = torch.empty((n, k), dtype=torch.int64, device=x.device) if I is None else I
            assert tuple(D.shape) == (n, k) and tuple(I.shape) == (n, k)
            _lib.check(self._L.bfk_search_sel(self._h, ctypes.c_void_p(xq.data_ptr()), n, k,
                                               ctypes.c_void_p(D.data_ptr()), ctypes.c_void_p(I.data_ptr()), 1,
                                               _torch_stream(self.device), _ALGO[algo], int(id_offset), _np_ptr(bm), 0, stp))
            torch.cuda.current_stream(self.device).synchronize()  # `bm` is host memory borrowed for the call
            return D, I
        x = np.ascontiguousarray(x, dtype="float32")
        D = np.empty((n, k), dtype=np.float32) if D is None else D
        I = np.empty((n, k), dtype=np.int64) if I is None else I
        assert D.shape == (n, k) and I.shape == (n, k)
        _lib.check(self._L.bfk_search_sel(self._h, _np_ptr(x), n, k, _np_ptr(D), _np_ptr(I), 0, None, _ALGO[algo],
                                           int(id_offset), _np_ptr(bm), 0, stp))
        return D, I

    def range_search(self, x, thresh, *, params=None):
        """faiss index.range_search(x, radius) -> (lims, D, I): all rows with squared L2 distance < radius
        (inner product > radius), grouped per query by lims[i]:lims[i+1]; rows of a query come in id order."""
        n, d = x.shape
        assert d == self.d
        if _is_torch(x):
            x = x.detach().to("cpu").numpy()
        x = np.ascontiguousarray(x, dtype="float32")
        lims = np.zeros(n + 1, dtype=np.int64)
        res = ctypes.c_void_p()
        _lib.check(self._L.bfk_range_search(self._h, _np_ptr(x), n, float(thresh), 0, None, _np_ptr(lims),
                                             ctypes.byref(res)))
        try:
            total = int(self._L.bfk_range_result_size(res))
            D = np.empty(total, dtype=np.float32)
            I = np.empty(total, dtype=np.int64)
            if total:
                _lib.check(self._L.bfk_range_result_copy(res, _np_ptr(D), _np_ptr(I), 0))
        finally:
            self._L.bfk_range_result_free(res)
        sel = getattr(params, "sel", None) if params is not None else None
        keep = None
        if sel is not None and total:
            bm = sel.to_bitmap(self.ntotal)
            keep = ((bm[I >> 3] >> (I & 7).astype(np.uint8)) & 1).astype(bool)
        # the kernel appends in scheduling order: make each query's rows deterministic (by id)
        out_l = np.zeros(n + 1, dtype=np.int64)
        Ds, Is = [], []
        for i in range(n):
            a, b = int(lims[i]), int(lims[i + 1])
            Di, Ii = D[a:b], I[a:b]
            if keep is not None:
                Di, Ii = Di[keep[a:b]], Ii[keep[a:b]]
            o = np.argsort(Ii, kind="stable")
            Ds.append(Di[o])
            Is.append(Ii[o])
            out_l[i + 1] = out_l[i] + Ii.size
        D = np.concatenate(Ds) if Ds else D
        I = np.concatenate(Is) if Is else I
        return out_l, D, I

    def finish(self) -> int:
        """Settle a `search(..., defer=True)`: wait, re-run uncertified queries, return rows rewritten.

        Extension (no faiss counterpart): with defer=True a CUDA-tensor search only enqueues work, so the
        caller can queue more GPU work behind it; (D, I) are final once finish() has returned.
        """
        n = ctypes.c_int64(0)
        _lib.check(self._L.bfk_search_finish(self._h, _torch_stream(self.device), ctypes.byref(n)))
        return int(n.value)

    def profile_begin(self) -> None:
        """Extension (measurement aid): start stamping every tensor-scan launch with the device clock."""
        _lib.check(self._L.bfk_profile_begin(self._h))

    def profile_end(self, detail: bool = False):
        """-> (summed scan-kernel milliseconds, number of scan launches) since profile_begin; waits for the GPU.
        detail=True -> dict with scan_ms, pass_ms (query preparation to re-rank end), idle_ms (device idle
        between consecutive passes: the host round trip) and n."""
        ms, ps, idle, n = ctypes.c_double(0.0), ctypes.c_double(0.0), ctypes.c_double(0.0), ctypes.c_int64(0)
        _lib.check(self._L.bfk_profile_end(self._h, ctypes.byref(ms), ctypes.byref(ps), ctypes.byref(idle), ctypes.byref(n)))
        if detail:
            return {"scan_ms": float(ms.value), "pass_ms": float(ps.value), "idle_ms": float(idle.value), "n": int(n.value)}
        return float(ms.value), int(n.value)

    def profile_stamps(self, max_passes: int = 65536) -> np.ndarray:
        """uint64 [n, 4] device-clock stamps (ns) of the passes since profile_begin: scan start, scan end,
        preparation start, re-rank end."""
        out = np.zeros((max_passes, 4), dtype=np.uint64)
        n = int(self._L.bfk_profile_stamps(self._h, _np_ptr(out), max_passes))
        if n < 0:
            raise RuntimeError("profile_stamps failed")
        return out[:n]

    def reconstruct(self, key, x=None):
        key = int(key)
        if x is None:
            x = np.empty(self.d, dtype=np.float32)
        else:
            assert x.shape == (self.d,)
        _lib.check(self._L.bfk_reconstruct(self._h, key, 1, _np_ptr(x), 0))
        return x

    def reconstruct_n(self, n0=0, ni=-1, x=None):
        if ni == -1:
            ni = self.ntotal
        if x is None:
            x = np.empty((ni, self.d), dtype=np.float32)
        else:
            assert x.shape == (ni, self.d)
        _lib.check(self._L.bfk_reconstruct(self._h, int(n0), int(ni), _np_ptr(x), 0))
        return x


class IndexFlatL2(IndexFlat):
    """faiss.IndexFlatL2 -- what semantic-indexer/indexer.py:39,104 constructs."""

    def __init__(self, d: int, **kw):
        super().__init__(d, METRIC_L2, **kw)


class IndexFlatIP(IndexFlat):
    """faiss.IndexFlatIP (BASELINE config 3; LangChain's MAX_INNER_PRODUCT strategy)."""

    def __init__(self, d: int, **kw):
        super().__init__(d, METRIC_INNER_PRODUCT, **kw)


def _wrap(handle, L) -> IndexFlat:
    metric = L.bfk_metric(handle)
    cls = IndexFlatL2 if metric == METRIC_L2 else IndexFlatIP
    obj = cls.__new__(cls)
    IndexFlat.__init__(obj, 0, _handle=handle)
    return obj


def write_index(index: IndexFlat, fname) -> None:
    """faiss.write_index -- indexer.py:27. Byte-identical IxF2 / IxFI file."""
    _lib.check(index._L.bfk_write(index._h, os.fsencode(fname)))


def write_index_append(index: IndexFlat, fname, rows_on_disk: int) -> None:
    """Extension of faiss.write_index for indexer.py:26-30 (a full rewrite per message): `fname` already holds
    the first `rows_on_disk` rows of `index`; only the rows after them are copied off the GPU and appended, then
    the header is updated in one write.  The file ends up byte-identical to write_index(index, fname).
    Raises RuntimeError when the file does not match (then write the whole file)."""
    _lib.check(index._L.bfk_write_append(index._h, os.fsencode(fname), int(rows_on_disk)))


def read_index(fname, *, device: int | None = None, storage="f32") -> IndexFlat:
    """faiss.read_index -- indexer.py:99, llm-qa/main.py:35."""
    L = _lib.load()
    h = ctypes.c_void_p()
    dev = _default_device() if device is None else int(device)
    _lib.check(L.bfk_read(os.fsencode(fname), dev, _STORAGE[storage], ctypes.byref(h)))
    return _wrap(h, L)


def serialize_index(index: IndexFlat) -> np.ndarray:
    n = int(index._L.bfk_serialized_size(index._h))
    buf = np.empty(n, dtype=np.uint8)
    _lib.check(index._L.bfk_serialize(index._h, _np_ptr(buf), n))
    return buf


def deserialize_index(data, *, device: int | None = None, storage="f32") -> IndexFlat:
    L = _lib.load()
    buf = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if isinstance(data, (bytes, bytearray)) else data,
                               dtype=np.uint8)
    h = ctypes.c_void_p()
    dev = _default_device() if device is None else int(device)
    _lib.check(L.bfk_deserialize(_np_ptr(buf), buf.size, dev, _STORAGE[storage], ctypes.byref(h)))
    return _wrap(h, L)


def normalize_L2(x: np.ndarray, *, device: int | None = None) -> None:
    """faiss.normalize_L2: in-place row normalisation of a float32 C-contiguous matrix."""
    assert x.dtype == np.float32 and x.flags.c_contiguous and x.ndim == 2
    dev = _default_device() if device is None else int(device)
    _lib.check(_lib.load().bfk_normalize_l2(_np_ptr(x), x.shape[0], x.shape[1], 0, dev, None))


def get_num_gpus() -> int:
    return int(_lib.load().bfk_device_count())

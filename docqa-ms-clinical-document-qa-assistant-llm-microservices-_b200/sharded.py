"""Row-sharded flat index: one process per GPU, NCCL all-gather of per-GPU top-k.

BASELINE config 5 (100M x 384 corpus over 2/4/8 B200).  The reference itself is a single
CPU process (semantic-indexer/indexer.py holds one faiss.IndexFlatL2; llm-qa/main.py:35
reads one file), so this class has no reference counterpart beyond the IndexFlat surface
it keeps: `add / search / ntotal` with the same (D, I) layout.

Data path per search:
    queries replicated on every rank
      -> local exact search on this rank's rows, ids already global (id_offset)
      -> ONE all-gather of the packed [nq, k] (D, I) lists  (torch.distributed, NCCL/NVLink)
      -> K6 merge kernel (bfk_merge_topk) on every rank
No other collective touches the data path; corpus rows never move between GPUs.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from ._lib import METRIC_INNER_PRODUCT, METRIC_L2


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous row range [lo, hi) of rank `rank` when n rows are split over `world`."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _cuda_merge(Dg, Ig, k: int, metric: int):
    """K6 through the C ABI: Dg/Ig are CUDA tensors [W, nq, k]."""
    import torch

    W, nq, _ = Dg.shape
    dev = Dg.device.index
    D = torch.empty((nq, k), dtype=torch.float32, device=Dg.device)
    I = torch.empty((nq, k), dtype=torch.int64, device=Dg.device)
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(_lib.load().bfk_merge_topk(ctypes.c_void_p(Dg.data_ptr()), ctypes.c_void_p(Ig.data_ptr()), W, nq, k,
                                           metric, ctypes.c_void_p(D.data_ptr()), ctypes.c_void_p(I.data_ptr()), dev,
                                           st))
    return D, I


class ShardedIndexFlat:
    """IndexFlat whose rows are split row-wise across the ranks of a process group."""

    def __init__(self, d: int, metric: int = METRIC_L2, *, group=None, device: int | None = None,
                 storage="f32", local_index=None, merge_fn=None):
        import torch.distributed as dist

        self._dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.d = int(d)
        self.metric_type = int(metric)
        self.is_trained = True
        if local_index is None:
            from .flat import IndexFlat

            local_index = IndexFlat(d, metric, device=device, storage=storage)
        self.local = local_index
        self._merge = merge_fn
        self._redo_check = True
        # piecewise-linear local -> global id map: local rows [ls, ls+n) are global [gs, gs+n)
        self._seg_local: list[int] = []
        self._seg_global: list[int] = []
        self._ntotal = 0

    # -- bookkeeping ------------------------------------------------------------------
    @property
    def ntotal(self) -> int:
        return self._ntotal

    def _record(self, n_local: int, global_start: int, n_global: int) -> None:
        if n_local:
            self._seg_local.append(self.local.ntotal - n_local)
            self._seg_global.append(global_start)
        self._n_records = getattr(self, "_n_records", 0) + 1  # the same on every rank, unlike the segment list
        self._ntotal += n_global

    # -- mutation -----------------------------------------------------------------------
    def add(self, x) -> None:
        """Replicated input: every rank passes the same [n, d]; each keeps its contiguous slice."""
        n, d = x.shape
        assert d == self.d
        lo, hi = shard_range(n, self.rank, self.world)
        if hi > lo:
            self.local.add(x[lo:hi])
        self._record(hi - lo, self._ntotal + lo, n)

    def add_local(self, x_local, global_start: int, n_global: int) -> None:
        """Rank-local rows whose global ids are [global_start, global_start + len(x_local))."""
        n, d = x_local.shape
        assert d == self.d
        if n:
            self.local.add(x_local)
        self._record(n, int(global_start), int(n_global))

    def reset(self) -> None:
        self.local.reset()
        self._seg_local, self._seg_global, self._ntotal = [], [], 0
        self._seg_tensors = None
        self._n_records = 0

    # -- queries ------------------------------------------------------------------------
    def _to_global(self, I):
        """Map local row ids to global ids (device or host tensor, -1 stays -1)."""
        import torch

        if len(self._seg_local) == 1:
            off = self._seg_global[0] - self._seg_local[0]
            return torch.where(I >= 0, I + off, I)
        if not self._seg_local:
            return I
        key = (str(I.device), len(self._seg_local))
        cached = getattr(self, "_seg_tensors", None)
        if cached is None or cached[0] != key:
            cached = (key, torch.tensor(self._seg_local, dtype=torch.int64, device=I.device),
                      torch.tensor(self._seg_global, dtype=torch.int64, device=I.device))
            self._seg_tensors = cached
        starts, gstarts = cached[1], cached[2]
        seg = torch.bucketize(I.clamp(min=0), starts, right=True) - 1
        out = gstarts[seg] + (I - starts[seg])
        return torch.where(I >= 0, out, I)

    def _exchange(self, nk, device):
        """The NVLink peer-memory exchange (K6x), created collectively on first use / when it must grow."""
        import torch

        L = _lib.load()
        if getattr(self, "_xh", None) is not None and L.bfk_xchg_capacity(self._xh) >= nk:
            return self._xh
        if getattr(self, "_xh", None) is not None:
            torch.cuda.synchronize(device)
            self._dist.barrier(group=self.group)
            L.bfk_xchg_free(self._xh)
            self._xh = None
        cap = max(int(nk) * 2, 1 << 16)
        h = ctypes.c_void_p()
        mine = np.zeros(64, dtype=np.uint8)
        _lib.check(L.bfk_xchg_create(device.index, self.world, self.rank, cap, ctypes.byref(h),
                                      ctypes.c_void_p(mine.ctypes.data)))
        # 64-byte CUDA IPC handles travel as 8 int64 words through the existing process group
        mine_t = torch.from_numpy(mine.view(np.int64).copy()).to(device)
        all_t = torch.empty((self.world, 8), dtype=torch.int64, device=device)
        self._dist.all_gather_into_tensor(all_t, mine_t, group=self.group)
        allh = np.ascontiguousarray(all_t.cpu().numpy()).view(np.uint8)
        _lib.check(L.bfk_xchg_connect(h, ctypes.c_void_p(allh.ctypes.data)))
        self._dist.barrier(group=self.group)
        self._xh = h
        return h

    def __del__(self):
        h, self._xh = getattr(self, "_xh", None), None
        if h:
            try:
                _lib.load().bfk_xchg_free(h)
            except Exception:
                pass

    def _gpu_buffers(self, n, k, device):
        """Per (n, k): one byte buffer holding this rank's D then I, the gather target, and the outputs."""
        import torch

        key = (n, k)
        buf = self._bufs.get(key) if hasattr(self, "_bufs") else None
        if buf is None:
            if not hasattr(self, "_bufs"):
                self._bufs = {}
            # int64 words: [ceil(n*k/2) words holding the fp32 distances][n*k words holding the ids]
            wd = (n * k + 1) // 2
            nwords = wd + n * k
            mine = torch.empty(nwords, dtype=torch.int64, device=device)
            allb = torch.empty((self.world, nwords), dtype=torch.int64, device=device)
            buf = {
                "mine": mine, "all": allb, "off_i": wd * 8, "nbytes": nwords * 8,
                "D": mine[:wd].view(torch.float32)[: n * k].view(n, k),
                "I": mine[wd:].view(n, k),
            }
            if len(self._bufs) > 16:
                self._bufs.clear()
            self._bufs[key] = buf
        return buf

    def search(self, x, k, *, algo=None):
        import torch

        n, d = x.shape
        assert d == self.d
        assert k > 0
        k = int(k)
        is_torch = type(x).__module__.split(".")[0] == "torch"
        # which exchange protocol runs must be the same decision on every rank: count add calls, not the local
        # segments (a rank that received no rows of a batch has fewer)
        single = getattr(self, "_n_records", 0) <= 1 and len(self._seg_local) <= 1
        one_add = getattr(self, "_n_records", 0) <= 1
        kw = {}
        if algo is not None:
            kw["algo"] = algo
        on_gpu = hasattr(self.local, "_h")
        if on_gpu and not one_add and self._merge is None and self.world > 1:
            # several id segments (incremental replicated adds): local ids are translated on the device, then
            # the same peer-memory exchange + merge; the local search is settled first (its rewrites would
            # otherwise land after the translation)
            dev = torch.device("cuda", self.local.device)
            b = self._gpu_buffers(n, k, dev)
            xq = x if is_torch else torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(dev)
            L = _lib.load()
            xh = self._exchange(n * k, dev)
            self.local.search(xq, k, D=b["D"], I=b["I"], **kw)
            b["I"].copy_(self._to_global(b["I"]))
            st = ctypes.c_void_p(torch.cuda.current_stream(dev.index).cuda_stream)
            Dout = torch.empty((n, k), dtype=torch.float32, device=dev)
            Iout = torch.empty((n, k), dtype=torch.int64, device=dev)
            _lib.check(L.bfk_xchg_merge(xh, ctypes.c_void_p(b["D"].data_ptr()), ctypes.c_void_p(b["I"].data_ptr()), n, k,
                                         self.metric_type, None, ctypes.c_void_p(Dout.data_ptr()),
                                         ctypes.c_void_p(Iout.data_ptr()), st))
            torch.cuda.current_stream(dev.index).synchronize()
            if is_torch:
                return Dout, Iout
            return Dout.cpu().numpy(), Iout.cpu().numpy()
        if on_gpu and one_add and self._merge is None and self.world > 1:
            # fast path: the kernels write (D, I) with global ids straight into the send buffer, ONE
            # all-gather of raw bytes, and the merge kernel reads the gather buffer in place
            dev = torch.device("cuda", self.local.device)
            b = self._gpu_buffers(n, k, dev)
            xq = x if is_torch else torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(dev)
            # the local search is only ENQUEUED (defer); the peer-memory exchange + merge kernels queue up
            # behind it, so the host never waits for the scan before launching them, and no host-launched
            # collective sits on the critical path
            L = _lib.load()
            xh = self._exchange(n * k, dev)
            off = self._seg_global[0] - self._seg_local[0] if self._seg_local else 0
            self.local.search(xq, k, D=b["D"], I=b["I"], id_offset=off, defer=True, **kw)
            st = ctypes.c_void_p(torch.cuda.current_stream(dev.index).cuda_stream)
            Dout = torch.empty((n, k), dtype=torch.float32, device=dev)
            Iout = torch.empty((n, k), dtype=torch.int64, device=dev)
            args = (ctypes.c_void_p(b["D"].data_ptr()), ctypes.c_void_p(b["I"].data_ptr()), n, k, self.metric_type)
            outs = (ctypes.c_void_p(Dout.data_ptr()), ctypes.c_void_p(Iout.data_ptr()), st)
            _lib.check(L.bfk_xchg_merge(xh, *args, L.bfk_search_pending_nflag_ptr(self.local._h), *outs))
            self.local.finish()  # waits for the stream when a tensor pass was pending; re-runs uncertified queries
            torch.cuda.current_stream(dev.index).synchronize()
            if L.bfk_xchg_last_max_nflag(xh) > 0:
                # some rank had to rewrite answers: every rank saw the same count, so all repeat the exchange
                _lib.check(L.bfk_xchg_merge(xh, *args, None, *outs))
                torch.cuda.current_stream(dev.index).synchronize()
            if is_torch:
                return Dout, Iout
            return Dout.cpu().numpy(), Iout.cpu().numpy()
        if single and on_gpu and self._seg_local:
            # ids come back global straight from the kernel's epilogue
            kw["id_offset"] = self._seg_global[0] - self._seg_local[0]
        Dl, Il = self.local.search(x, k, **kw)
        if isinstance(Dl, np.ndarray):
            Dl, Il = torch.from_numpy(np.ascontiguousarray(Dl)), torch.from_numpy(np.ascontiguousarray(Il))
        if on_gpu and not Dl.is_cuda:
            dev = torch.device("cuda", self.local.device)
            Dl, Il = Dl.to(dev), Il.to(dev)
        if "id_offset" not in kw:
            Il = self._to_global(Il)
        if self.world == 1:
            D, I = Dl, Il
        else:
            # pack D (fp32 bits) and I into one int64 payload: a single all-gather per search
            payload = torch.stack([Dl.contiguous().view(torch.int32).to(torch.int64), Il.contiguous()], dim=0)
            flat = torch.empty((self.world * 2, n, k), dtype=torch.int64, device=payload.device)
            self._dist.all_gather_into_tensor(flat, payload, group=self.group)
            gathered = flat.view(self.world, 2, n, k)
            Dg = gathered[:, 0].to(torch.int32).view(torch.float32).contiguous()
            Ig = gathered[:, 1].contiguous()
            if self._merge is not None:
                D, I = self._merge(Dg, Ig, k, self.metric_type)
            elif Dg.is_cuda:
                D, I = _cuda_merge(Dg, Ig, k, self.metric_type)
            else:
                raise RuntimeError("ShardedIndexFlat: merging needs the CUDA library (no CPU fallback)")
        if is_torch:
            return D, I
        return D.cpu().numpy(), I.cpu().numpy()

"""Row-sharded flat index: one process per GPU, per-GPU top-k lists exchanged and merged.

BASELINE config 5 (100M x 384 corpus over 2/4/8 B200).  The reference itself is a single
CPU process (semantic-indexer/indexer.py holds one faiss.IndexFlatL2; llm-qa/main.py:35
reads one file), so this class keeps the IndexFlat surface those call sites use --
`add / search / ntotal / reconstruct(_n) / remove_ids / reset / range_search`, module-level
`write_index / read_index` equivalents as methods -- on an index whose rows live on W GPUs.
Every method is COLLECTIVE: all ranks call it with the same arguments, all get the same result.

Placement (which rank holds global row g; ids are positional like faiss's):
  * "block_cyclic" (default; `add`): blocks of `block_rows` = 65536 consecutive rows go round the
    ranks, so a stream of small adds (the reference appends ONE row per call, indexer.py:41) fills
    all GPUs alike and the id map stays arithmetic in both directions:
        owner(g) = (g // B) % W        local(g) = (g // (B*W)) * B + g % B
        global(l, r) = ((l // B) * W + r) * B + l % B
  * "contiguous" (`add_local` / `set_contiguous_shard`): rank r holds one contiguous range, declared
    by the caller (bulk loads, the benchmark); ids come out of the kernels global via `id_offset`.

Data path per search:
    queries replicated on every rank
      -> local exact search on this rank's rows
      -> the [nq, k] (D, I) lists of every rank reach every rank: over NVLink peer memory, fused with the
         merge (K6x, csrc/bfk_xchg.cu) -- or ONE NCCL/gloo all-gather where peers cannot map each other
      -> K6 merge kernel on every rank
No other collective touches the data path; corpus rows never move between GPUs.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from . import _lib
from ._lib import METRIC_INNER_PRODUCT, METRIC_L2

BLOCK_ROWS = 65536
_HEADER_BYTES = 45


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


def _faiss_header(d: int, n: int, metric: int) -> bytes:
    """The 45-byte IxF2 / IxFI header (same bytes as csrc/bfk_index.cu put_header and faiss's writer)."""
    h = bytearray(_HEADER_BYTES)
    h[0:4] = b"IxF2" if metric == METRIC_L2 else b"IxFI"
    h[4:8] = np.int32(d).tobytes()
    h[8:16] = np.int64(n).tobytes()
    h[16:24] = np.int64(1 << 20).tobytes()
    h[24:32] = np.int64(1 << 20).tobytes()
    h[32] = 1
    h[33:37] = np.int32(metric).tobytes()
    h[37:45] = np.uint64(n * d).tobytes()
    return bytes(h)


def _parse_faiss_header(raw: bytes):
    if len(raw) < _HEADER_BYTES or raw[:4] not in (b"IxF2", b"IxFI"):
        raise RuntimeError(f"Index type {raw[:4]!r} not recognized (only flat L2/IP)")
    d = int(np.frombuffer(raw[4:8], dtype=np.int32)[0])
    n = int(np.frombuffer(raw[8:16], dtype=np.int64)[0])
    count = int(np.frombuffer(raw[37:45], dtype=np.uint64)[0])
    if d <= 0 or n < 0 or count != n * d:
        raise RuntimeError("flat index: corrupt header")
    return d, n, METRIC_L2 if raw[:4] == b"IxF2" else METRIC_INNER_PRODUCT


class ShardedIndexFlat:
    """IndexFlat whose rows are split row-wise across the ranks of a process group."""

    def __init__(self, d: int, metric: int = METRIC_L2, *, group=None, device: int | None = None,
                 storage="f32", local_index=None, merge_fn=None, block_rows: int = BLOCK_ROWS):
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
        self.block_rows = int(block_rows)
        assert self.block_rows > 0
        self.placement = None     # "block_cyclic" / "contiguous", fixed by the first add
        self._global_start = 0    # contiguous: global id of local row 0
        self._ntotal = 0
        self._xh = None
        self._pin = {}

    # -- bookkeeping ------------------------------------------------------------------
    @property
    def ntotal(self) -> int:
        return self._ntotal

    @property
    def on_gpu(self) -> bool:
        return hasattr(self.local, "_h")

    def _set_placement(self, kind: str) -> None:
        if self.placement is None:
            self.placement = kind
        elif self.placement != kind:
            raise RuntimeError(f"this index is {self.placement}; `add` (block-cyclic) and `add_local` (one contiguous "
                               "shard per rank) cannot be mixed")

    # id arithmetic (numpy or torch integer arrays alike)
    def _owned_in(self, g0: int, g1: int):
        """Global rows of [g0, g1) this rank owns, as a list of (start, stop) runs in increasing order."""
        if g1 <= g0:
            return []
        if self.placement == "contiguous":
            lo, hi = self._global_start, self._global_start + int(self.local.ntotal)
            a, b = max(g0, lo), min(g1, hi)
            return [(a, b)] if b > a else []
        B, W, r = self.block_rows, self.world, self.rank
        runs = []
        b = g0 // B
        # first block of this rank at or after b
        b += (r - b) % W
        while b * B < g1:
            a, e = max(g0, b * B), min(g1, (b + 1) * B)
            if e > a:
                runs.append((a, e))
            b += W
        return runs

    def _local_of(self, g: int) -> int:
        if self.placement == "contiguous":
            return g - self._global_start
        B, W = self.block_rows, self.world
        return (g // (B * W)) * B + g % B

    def _to_global(self, I):
        """Local row ids -> global ids (numpy or torch, -1 stays -1)."""
        if self.placement == "contiguous" or self.placement is None:
            off = self._global_start
            if off == 0:
                return I
            return I + off * (I >= 0)
        B, W, r = self.block_rows, self.world, self.rank
        g = ((I // B) * W + r) * B + I % B
        if isinstance(I, np.ndarray):
            return np.where(I >= 0, g, I)
        import torch

        return torch.where(I >= 0, g, I)

    def _local_global_ids(self) -> np.ndarray:
        """Global id of every local row, in local order."""
        n = int(self.local.ntotal)
        return np.asarray(self._to_global(np.arange(n, dtype=np.int64)))

    # -- mutation -----------------------------------------------------------------------
    def add(self, x) -> None:
        """index.add(x) -- indexer.py:41.  Replicated input: every rank passes the same [n, d] and keeps the
        rows of its own blocks; the new rows get the next n global ids."""
        n, d = x.shape
        assert d == self.d
        self._set_placement("block_cyclic")
        g0 = self._ntotal
        for a, b in self._owned_in(g0, g0 + n):
            self.local.add(x[a - g0 : b - g0])
        self._ntotal += n

    def add_local(self, x_local, global_start: int, n_global: int) -> None:
        """Bulk load: this rank's rows, whose global ids are [global_start, global_start + len(x_local)); the
        ranks' ranges tile [ntotal, ntotal + n_global) in rank order."""
        n, d = x_local.shape
        assert d == self.d
        if n:
            self.local.add(x_local)
        self.set_contiguous_shard(int(global_start) - (int(self.local.ntotal) - n), int(n_global))

    def set_contiguous_shard(self, global_start: int, n_global: int) -> None:
        """Declares that local rows [0, local.ntotal) -- however they were added -- are global rows
        [global_start, global_start + local.ntotal) of an index that now holds n_global more rows."""
        self._set_placement("contiguous")
        if self._ntotal and int(global_start) != self._global_start:
            raise RuntimeError("a contiguous shard has one fixed global range per rank")
        self._global_start = int(global_start)
        self._ntotal += int(n_global)

    def reset(self) -> None:
        self.local.reset()
        self._ntotal, self._global_start, self.placement = 0, 0, None

    def remove_ids(self, ids) -> int:
        """faiss remove_ids: the surviving rows keep their order and are renumbered 0..ntotal-1.  Renumbering moves
        rows to other blocks, so the survivors are re-dealt to their new owners (one all-to-all of rows)."""
        import torch

        ids = np.unique(np.asarray(ids, dtype=np.int64).reshape(-1))
        ids = ids[(ids >= 0) & (ids < self._ntotal)]
        if ids.size == 0:
            return 0
        if self.placement == "contiguous":
            raise RuntimeError("remove_ids needs the block-cyclic placement (build the index with `add`)")
        g = self._local_global_ids()
        keep = ~np.isin(g, ids)
        new_g = g[keep] - np.searchsorted(ids, g[keep])
        rows = np.ascontiguousarray(self.local.reconstruct_n(0, int(self.local.ntotal))[keep], dtype=np.float32) \
            if g.size else np.zeros((0, self.d), np.float32)
        B, W = self.block_rows, self.world
        dest = (new_g // B) % W
        order = np.argsort(dest, kind="stable")  # new_g is increasing, so each destination's rows stay ordered
        send_counts = np.bincount(dest, minlength=W).astype(np.int64)
        dev = self._comm_device()
        sc = torch.from_numpy(send_counts).to(dev)
        rc = torch.empty(W, dtype=torch.int64, device=dev)
        self._dist.all_to_all_single(rc, sc, group=self.group)
        recv_counts = rc.cpu().numpy()
        send_rows = torch.from_numpy(rows[order]).to(dev)
        send_ids = torch.from_numpy(new_g[order]).to(dev)
        m = int(recv_counts.sum())
        recv_rows = torch.empty((m, self.d), dtype=torch.float32, device=dev)
        recv_ids = torch.empty(m, dtype=torch.int64, device=dev)
        self._dist.all_to_all_single(recv_rows, send_rows, (recv_counts).tolist(), (send_counts).tolist(), group=self.group)
        self._dist.all_to_all_single(recv_ids, send_ids, recv_counts.tolist(), send_counts.tolist(), group=self.group)
        o = torch.argsort(recv_ids)
        new_rows = recv_rows[o].cpu().numpy()
        self.local.reset()
        if m:
            self.local.add(new_rows)
        self._ntotal -= int(ids.size)
        # local order == increasing new global id == the arithmetic placement
        assert m == sum(b - a for a, b in self._owned_in(0, self._ntotal))
        return int(ids.size)

    def _comm_device(self):
        import torch

        backend = self._dist.get_backend(self.group)
        return torch.device("cuda", self.local.device) if ("nccl" in str(backend) and self.on_gpu) else torch.device("cpu")

    # -- reading rows back ------------------------------------------------------------------
    def reconstruct_n(self, n0: int = 0, ni: int = -1, x=None):
        """Rows [n0, n0 + ni) in global order on every rank (each rank contributes its own, one all-reduce)."""
        import torch

        if ni == -1:
            ni = self._ntotal - n0
        assert 0 <= n0 and n0 + ni <= self._ntotal
        out = np.zeros((ni, self.d), dtype=np.float32)
        for a, b in self._owned_in(n0, n0 + ni):
            out[a - n0 : b - n0] = self.local.reconstruct_n(self._local_of(a), b - a)
        if self.world > 1 and ni:
            t = torch.from_numpy(out).to(self._comm_device())
            self._dist.all_reduce(t, group=self.group)  # exactly one rank holds each row, the others add zeros
            out = t.cpu().numpy()
        if x is not None:
            x[...] = out
            return x
        return out

    def reconstruct(self, key: int, x=None):
        row = self.reconstruct_n(int(key), 1)[0]
        if x is not None:
            x[...] = row
            return x
        return row

    def write_index(self, fname) -> None:
        """faiss.write_index(index, fname) -- indexer.py:27 -- for the sharded index: ONE file in faiss's format,
        byte-identical to what a single-GPU index holding the same rows writes.  Every rank writes its own rows
        at their offsets (the ranks share the node's file system); rank 0 writes the header last, so the file
        only becomes a valid index once all rows are in place."""
        fname = os.fspath(fname)
        total = _HEADER_BYTES + 4 * self.d * self._ntotal
        if self.rank == 0:
            with open(fname, "wb") as f:
                f.truncate(total)
        self._barrier()
        chunk = max(1, (64 << 20) // (4 * self.d))
        with open(fname, "r+b") as f:
            for a, b in self._owned_in(0, self._ntotal):
                for c0 in range(a, b, chunk):
                    c1 = min(b, c0 + chunk)
                    rows = np.ascontiguousarray(self.local.reconstruct_n(self._local_of(c0), c1 - c0), dtype=np.float32)
                    os.pwrite(f.fileno(), rows.tobytes(), _HEADER_BYTES + 4 * self.d * c0)
            f.flush()
            os.fsync(f.fileno())
        self._barrier()
        if self.rank == 0:
            with open(fname, "r+b") as f:
                os.pwrite(f.fileno(), _faiss_header(self.d, self._ntotal, self.metric_type), 0)
                f.flush()
                os.fsync(f.fileno())
        self._barrier()

    @classmethod
    def read_index(cls, fname, *, group=None, device: int | None = None, storage="f32", local_index_factory=None,
                   merge_fn=None, placement: str = "block_cyclic", block_rows: int = BLOCK_ROWS):
        """faiss.read_index(fname) -- indexer.py:99, llm-qa/main.py:35 -- onto W GPUs: every rank reads the header
        and only its own rows of the file (its blocks, or its `shard_range` for placement="contiguous")."""
        fname = os.fspath(fname)
        with open(fname, "rb") as f:
            d, n, metric = _parse_faiss_header(f.read(_HEADER_BYTES))
        local = local_index_factory(d, metric) if local_index_factory else None
        self = cls(d, metric, group=group, device=device, storage=storage, local_index=local, merge_fn=merge_fn,
                   block_rows=block_rows)
        chunk = max(1, (64 << 20) // (4 * d))

        def rows_of(a, b):
            for c0 in range(a, b, chunk):
                c1 = min(b, c0 + chunk)
                yield np.fromfile(fname, dtype=np.float32, count=(c1 - c0) * d, offset=_HEADER_BYTES + 4 * d * c0).reshape(-1, d)

        if placement == "contiguous":
            lo, hi = shard_range(n, self.rank, self.world)
            if hasattr(self.local, "reserve"):
                self.local.reserve(max(hi - lo, 1))
            for rows in rows_of(lo, hi):
                self.local.add(rows)
            self.set_contiguous_shard(lo, n)
        else:
            self.placement = "block_cyclic"
            runs = self._owned_in(0, n)
            if hasattr(self.local, "reserve"):
                self.local.reserve(max(sum(b - a for a, b in runs), 1))
            for a, b in runs:
                for rows in rows_of(a, b):
                    self.local.add(rows)
            self._ntotal = n
        return self

    def _barrier(self) -> None:
        if self.world > 1:
            self._dist.barrier(group=self.group)

    # -- the peer-memory exchange ---------------------------------------------------------------
    def _exchange(self, nk, device):
        """The NVLink peer-memory exchange (K6x), created collectively on first use / when it must grow."""
        import torch

        L = _lib.load()
        if self._xh is not None and L.bfk_xchg_capacity(self._xh) >= nk:
            return self._xh
        if self._xh is not None:
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
        """Per (n, k): this rank's (D, I) send buffer, the merged outputs, and pinned host mirrors."""
        import torch

        key = (n, k)
        buf = self._pin.get(key)
        if buf is None:
            if len(self._pin) > 16:
                self._pin.clear()
            buf = {
                "D": torch.empty((n, k), dtype=torch.float32, device=device),
                "I": torch.empty((n, k), dtype=torch.int64, device=device),
                "Dout": torch.empty((n, k), dtype=torch.float32, device=device),  # host-side callers only
                "Iout": torch.empty((n, k), dtype=torch.int64, device=device),
                "q_dev": torch.empty((n, self.d), dtype=torch.float32, device=device),
                # host-side callers: queries go up and results come down through pinned memory, asynchronously,
                # with ONE wait at the end (a pageable .to(device) / .cpu() is a synchronous round trip each)
                "q_pin": torch.empty((n, self.d), dtype=torch.float32, pin_memory=True),
                "D_pin": torch.empty((n, k), dtype=torch.float32, pin_memory=True),
                "I_pin": torch.empty((n, k), dtype=torch.int64, pin_memory=True),
            }
            self._pin[key] = buf
        return buf

    # -- queries ------------------------------------------------------------------------
    def search(self, x, k, *, params=None, algo=None):
        import torch

        n, d = x.shape
        assert d == self.d
        assert k > 0
        k = int(k)
        is_torch = type(x).__module__.split(".")[0] == "torch"
        kw = {}
        if algo is not None:
            kw["algo"] = algo
        sel = getattr(params, "sel", None) if params is not None else None
        if self.on_gpu and self._merge is None and self.world > 1 and sel is None and n * k <= (1 << 22):
            return self._search_peer_exchange(x, n, k, is_torch, kw)

        # ---- general route: local search, ONE all-gather of the packed lists, merge ----
        if sel is not None:
            kw["params"] = self._local_params(sel)
        contiguous = self.placement != "block_cyclic"
        if contiguous and self.on_gpu and self._global_start:
            kw["id_offset"] = self._global_start  # ids come back global straight from the kernel's epilogue
        Dl, Il = self.local.search(x, k, **kw)
        if isinstance(Dl, np.ndarray):
            Dl, Il = torch.from_numpy(np.ascontiguousarray(Dl)), torch.from_numpy(np.ascontiguousarray(Il))
        if self.on_gpu and not Dl.is_cuda:
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

    def _search_peer_exchange(self, x, n, k, is_torch, kw):
        """Fast path: the local search is only ENQUEUED (defer); the kernels write (D, I) straight into the send
        buffer; the peer-memory exchange + merge kernels queue up behind it, so the host never waits for the scan
        before launching them and no host-launched collective sits on the critical path."""
        import torch

        dev = torch.device("cuda", self.local.device)
        b = self._gpu_buffers(n, k, dev)
        stream = torch.cuda.current_stream(dev.index)
        if is_torch:
            xq = x
        else:
            b["q_pin"].numpy()[...] = x
            xq = b["q_dev"]
            xq.copy_(b["q_pin"], non_blocking=True)
        L = _lib.load()
        xh = self._exchange(n * k, dev)
        block_cyclic = self.placement == "block_cyclic"
        _lib.check(L.bfk_xchg_set_id_map(xh, self.block_rows if block_cyclic else 0))
        off = 0 if block_cyclic else self._global_start
        self.local.search(xq, k, D=b["D"], I=b["I"], id_offset=off, defer=True, **kw)
        st = ctypes.c_void_p(stream.cuda_stream)
        args = (ctypes.c_void_p(b["D"].data_ptr()), ctypes.c_void_p(b["I"].data_ptr()), n, k, self.metric_type)
        if is_torch:  # the caller keeps these; host-side callers get copies of the persistent pair
            Dout = torch.empty((n, k), dtype=torch.float32, device=dev)
            Iout = torch.empty((n, k), dtype=torch.int64, device=dev)
        else:
            Dout, Iout = b["Dout"], b["Iout"]
        outs = (ctypes.c_void_p(Dout.data_ptr()), ctypes.c_void_p(Iout.data_ptr()), st)
        _lib.check(L.bfk_xchg_merge(xh, *args, L.bfk_search_pending_nflag_ptr(self.local._h), *outs))
        if not is_torch:
            b["D_pin"].copy_(Dout, non_blocking=True)
            b["I_pin"].copy_(Iout, non_blocking=True)
        self.local.finish()  # waits for the local pass when one was pending; re-runs uncertified queries
        stream.synchronize()
        status = L.bfk_xchg_last_max_nflag(xh)
        if status == -2:
            raise RuntimeError("sharded search: a rank's lists did not arrive within the exchange time limit "
                               "(BFK_XCHG_TIMEOUT_S); the process group is out of step")
        if status > 0:
            # some rank had to rewrite answers: every rank saw the same count, so all repeat the exchange
            _lib.check(L.bfk_xchg_merge(xh, *args, None, *outs))
            if not is_torch:
                b["D_pin"].copy_(Dout, non_blocking=True)
                b["I_pin"].copy_(Iout, non_blocking=True)
            stream.synchronize()
        if is_torch:
            return Dout, Iout
        return b["D_pin"].numpy().copy(), b["I_pin"].numpy().copy()

    def _local_params(self, sel):
        """faiss SearchParameters(sel=...) over GLOBAL ids -> the same restricted to this rank's rows, in local order."""
        from .idselect import IDSelectorBitmap, SearchParameters

        bm = np.asarray(sel.to_bitmap(self._ntotal), dtype=np.uint8)
        g = self._local_global_ids()
        bits = ((bm[g >> 3] >> (g & 7).astype(np.uint8)) & 1).astype(bool) if g.size else np.zeros(0, bool)
        return SearchParameters(sel=IDSelectorBitmap.from_mask(bits))

    def range_search(self, x, thresh, *, params=None):
        """faiss range_search -> (lims, D, I) with global ids; each query's rows in id order (a side path: the
        per-rank results travel as Python objects through all_gather_object)."""
        n, d = x.shape
        assert d == self.d
        kw = {"params": self._local_params(params.sel)} if (params is not None and getattr(params, "sel", None) is not None) else {}
        if type(x).__module__.split(".")[0] == "torch":
            x = x.detach().cpu().numpy()
        lims, D, I = self.local.range_search(np.ascontiguousarray(x, dtype=np.float32), thresh, **kw)
        I = np.asarray(self._to_global(np.asarray(I, dtype=np.int64)))
        parts = [None] * self.world
        if self.world > 1:
            self._dist.all_gather_object(parts, (np.asarray(lims), np.asarray(D), I), group=self.group)
        else:
            parts = [(np.asarray(lims), np.asarray(D), I)]
        out_l = np.zeros(n + 1, dtype=np.int64)
        Ds, Is = [], []
        for q in range(n):
            Dq = np.concatenate([p[1][p[0][q]:p[0][q + 1]] for p in parts]) if parts else np.zeros(0, np.float32)
            Iq = np.concatenate([p[2][p[0][q]:p[0][q + 1]] for p in parts]) if parts else np.zeros(0, np.int64)
            o = np.argsort(Iq, kind="stable")
            Ds.append(Dq[o])
            Is.append(Iq[o])
            out_l[q + 1] = out_l[q] + Iq.size
        return out_l, (np.concatenate(Ds) if Ds else np.zeros(0, np.float32)), (np.concatenate(Is) if Is else np.zeros(0, np.int64))

"""ORACLE (test infrastructure, not product code): faiss flat-index file format.

CPU restatement of what `faiss.write_index` / `faiss.read_index` do for an
`IndexFlatL2` / `IndexFlatIP`, i.e. the on-disk contract between
`semantic-indexer/indexer.py:27` (writer), `indexer.py:99` and
`llm-qa/main.py:35` (readers).  The algorithm lives in the third-party
`faiss-cpu` wheel (unpinned in `semantic-indexer/requirements.txt:2`); it is
absent from /root/reference and from this image, so the layout below is
restated from the published format and *pinned* against the reference's own
committed artefact `semantic-indexer/vector_store.faiss` (996 909 bytes),
whose bytes this module must reproduce exactly (tests/test_oracle_io.py).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may
import this module.

Layout (little endian):
    [0:4]   fourcc  "IxF2" (L2) | "IxFI" (inner product) | legacy "IxFl"
    [4:8]   int32   d
    [8:16]  int64   ntotal
    [16:24] int64   dummy 1<<20
    [24:32] int64   dummy 1<<20
    [32]    uint8   is_trained
    [33:37] int32   metric_type (0 = IP, 1 = L2)
    [37:41] float32 metric_arg, only when metric_type > 1
    [..+8]  uint64  count = ntotal*d (number of float32 that follow)
    [..]    count float32, row major
"""
from __future__ import annotations

import struct

import numpy as np

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1

_FOURCC = {METRIC_L2: b"IxF2", METRIC_INNER_PRODUCT: b"IxFI"}


def write_index_bytes(x: np.ndarray, metric: int) -> bytes:
    """Serialise an (ntotal, d) float32 matrix as faiss would."""
    x = np.ascontiguousarray(x, dtype="<f4")
    if x.ndim != 2:
        raise ValueError("x must be 2-D")
    n, d = x.shape
    if metric not in _FOURCC:
        raise ValueError("only METRIC_L2 / METRIC_INNER_PRODUCT flat indexes")
    head = _FOURCC[metric]
    head += struct.pack("<i", d)
    head += struct.pack("<q", n)
    head += struct.pack("<q", 1 << 20)
    head += struct.pack("<q", 1 << 20)
    head += struct.pack("<B", 1)
    head += struct.pack("<i", metric)
    head += struct.pack("<Q", n * d)
    return head + x.tobytes()


def read_index_bytes(buf: bytes):
    """Parse a flat index file -> (d, ntotal, metric, x[ntotal, d] float32)."""
    if len(buf) < 45:
        raise RuntimeError("file too short for a flat index header")
    fourcc = bytes(buf[0:4])
    if fourcc not in (b"IxF2", b"IxFI", b"IxFl"):
        raise RuntimeError(f"Index type {fourcc!r} not recognized")
    (d,) = struct.unpack_from("<i", buf, 4)
    (ntotal,) = struct.unpack_from("<q", buf, 8)
    (is_trained,) = struct.unpack_from("<B", buf, 32)
    (metric,) = struct.unpack_from("<i", buf, 33)
    off = 37
    if metric > 1:
        off += 4
    (count,) = struct.unpack_from("<Q", buf, off)
    off += 8
    if count != ntotal * d:
        raise RuntimeError("corrupt flat index: count != ntotal*d")
    if len(buf) < off + 4 * count:
        raise RuntimeError("corrupt flat index: truncated payload")
    x = np.frombuffer(buf, dtype="<f4", count=count, offset=off).reshape(ntotal, d)
    if fourcc == b"IxF2":
        metric = METRIC_L2
    elif fourcc == b"IxFI":
        metric = METRIC_INNER_PRODUCT
    return d, ntotal, metric, np.array(x, dtype=np.float32)

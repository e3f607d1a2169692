"""Masked mean-pool + L2-normalise on the GPU (kernel K1, csrc/bfk_pool.cu).

The tail of `SentenceTransformer('all-MiniLM-L6-v2').encode` that the reference reaches
at semantic-indexer/indexer.py:37 and llm-qa/main.py:25:
    e = sum_t tok[t]*m[t] / max(sum_t m[t], 1e-9);  v = e / max(||e||_2, 1e-12)
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from . import _lib
from ._lib import DT_BF16, DT_F16, DT_F32, DT_I32, DT_I64, DT_U8


def _tok_dtype(t) -> int:
    import torch

    return {torch.float32: DT_F32, torch.float16: DT_F16, torch.bfloat16: DT_BF16}[t.dtype]


def _mask_dtype(t) -> int:
    import torch

    return {torch.int64: DT_I64, torch.int32: DT_I32, torch.uint8: DT_U8, torch.bool: DT_U8}[t.dtype]


_NP_TOK = {np.dtype("float32"): DT_F32, np.dtype("float16"): DT_F16}
_NP_MASK = {np.dtype("int64"): DT_I64, np.dtype("int32"): DT_I32, np.dtype("uint8"): DT_U8, np.dtype("bool"): DT_U8}


def pool_normalize(token_embeddings, attention_mask, *, device: int | None = None):
    """[B, T, H] token states + [B, T] mask -> [B, H] float32 unit-norm sentence vectors.

    CUDA torch tensors stay on the device (result is a CUDA tensor); numpy arrays are
    copied to the GPU and back.  There is no CPU implementation.
    """
    L = _lib.load()
    if type(token_embeddings).__module__.split(".")[0] == "torch":
        import torch

        tok = token_embeddings.contiguous()
        mask = attention_mask.contiguous()
        B, T, H = tok.shape
        assert tuple(mask.shape) == (B, T)
        assert tok.is_cuda and mask.is_cuda
        dev = tok.device.index
        out = torch.empty((B, H), dtype=torch.float32, device=tok.device)
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(L.bfk_pool_normalize(ctypes.c_void_p(tok.data_ptr()), _tok_dtype(tok),
                                         ctypes.c_void_p(mask.data_ptr()), _mask_dtype(mask), B, T, H,
                                         ctypes.c_void_p(out.data_ptr()), 1, dev, st))
        return out
    tok = np.ascontiguousarray(token_embeddings)
    if tok.dtype not in _NP_TOK:
        tok = tok.astype(np.float32)
    mask = np.ascontiguousarray(attention_mask)
    if mask.dtype not in _NP_MASK:
        mask = mask.astype(np.int64)
    B, T, H = tok.shape
    assert mask.shape == (B, T)
    out = np.empty((B, H), dtype=np.float32)
    dev = int(os.environ.get("BFK_DEVICE", "0")) if device is None else int(device)
    _lib.check(L.bfk_pool_normalize(ctypes.c_void_p(tok.ctypes.data), _NP_TOK[tok.dtype],
                                     ctypes.c_void_p(mask.ctypes.data), _NP_MASK[mask.dtype], B, T, H,
                                     ctypes.c_void_p(out.ctypes.data), 0, dev, None))
    return out

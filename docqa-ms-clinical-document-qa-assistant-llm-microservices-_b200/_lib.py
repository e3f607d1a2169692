"""ctypes binding of libbfk.so (the C ABI declared in include/bfk.h).

There is no CPU fallback: if the shared library is missing the import of the compute
surface fails loudly, and every compute entry point fails without a CUDA device.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libbfk.so")

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1
STORE_F32 = 0
STORE_BF16 = 1
DT_F32, DT_F16, DT_BF16, DT_I64, DT_I32, DT_U8 = 0, 1, 2, 3, 4, 5
ALGO_AUTO, ALGO_EXACT_SIMT, ALGO_TENSOR = 0, 1, 2
ALGO_FLAG_DEFER = 0x100
MAX_K = 2048
MAX_D = 2048


class SearchStats(ctypes.Structure):
    _fields_ = [
        ("algo_used", ctypes.c_int32),
        ("n_launches", ctypes.c_int32),
        ("n_fallback", ctypes.c_int64),
        ("n_candidates", ctypes.c_int64),
        ("scan_ms", ctypes.c_float),
        ("reserved", ctypes.c_float),
    ]


# every symbol include/bfk.h declares; tests/test_cabi_symbols.py checks the .so exports all
EXPORTS = [
    "bfk_last_error", "bfk_version", "bfk_device_count", "bfk_create", "bfk_free", "bfk_dim",
    "bfk_metric", "bfk_storage", "bfk_device", "bfk_ntotal", "bfk_reserve", "bfk_add",
    "bfk_add_pooled", "bfk_pool_normalize", "bfk_search", "bfk_search_ex", "bfk_search_finish", "bfk_search_pending_nflag_ptr", "bfk_reset",
    "bfk_xchg_create", "bfk_xchg_connect", "bfk_xchg_capacity", "bfk_xchg_merge", "bfk_xchg_last_max_nflag",
    "bfk_xchg_free", "bfk_xchg_set_id_map",
    "bfk_reconstruct", "bfk_remove_ids", "bfk_write", "bfk_write_append", "bfk_read", "bfk_serialized_size",
    "bfk_serialize", "bfk_deserialize", "bfk_normalize_l2", "bfk_merge_topk", "bfk_merge_topk_strided", "bfk_search_noalloc",
    "bfk_enc_create", "bfk_enc_free", "bfk_enc_set_tensor", "bfk_enc_forward", "bfk_enc_encode", "bfk_enc_encode_add",
    "bfk_profile_begin", "bfk_profile_end", "bfk_profile_stamps",
    "bfk_search_sel", "bfk_range_search", "bfk_range_result_size", "bfk_range_result_copy", "bfk_range_result_free",
]

_lib = None


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise RuntimeError(
            f"{SO_PATH} is missing: build it with `python {os.path.join(_HERE, 'build.py')}` "
            "(libbfk has no CPU fallback)"
        )
    L = ctypes.CDLL(SO_PATH)
    vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    L.bfk_last_error.restype = ctypes.c_char_p
    L.bfk_last_error.argtypes = []
    L.bfk_version.restype = i32
    L.bfk_device_count.restype = i32
    L.bfk_create.argtypes = [i32, i32, i32, i32, ctypes.POINTER(vp)]
    L.bfk_free.argtypes = [vp]
    for name in ("bfk_dim", "bfk_metric", "bfk_storage", "bfk_device"):
        getattr(L, name).argtypes = [vp]
        getattr(L, name).restype = i32
    L.bfk_ntotal.argtypes = [vp]
    L.bfk_ntotal.restype = i64
    L.bfk_reserve.argtypes = [vp, i64]
    L.bfk_add.argtypes = [vp, vp, i64, i32, vp]
    L.bfk_add_pooled.argtypes = [vp, vp, i32, vp, i32, i64, i32, i32, vp]
    L.bfk_pool_normalize.argtypes = [vp, i32, vp, i32, i64, i32, i32, vp, i32, i32, vp]
    L.bfk_search.argtypes = [vp, vp, i64, i64, vp, vp, i32, vp]
    L.bfk_search_ex.argtypes = [vp, vp, i64, i64, vp, vp, i32, vp, i32, i64, ctypes.POINTER(SearchStats)]
    L.bfk_search_finish.argtypes = [vp, vp, ctypes.POINTER(i64)]
    L.bfk_profile_begin.argtypes = [vp]
    dbl_p = ctypes.POINTER(ctypes.c_double)
    L.bfk_profile_end.argtypes = [vp, dbl_p, dbl_p, dbl_p, ctypes.POINTER(i64)]
    L.bfk_profile_stamps.argtypes = [vp, vp, i64]
    L.bfk_profile_stamps.restype = i64
    L.bfk_search_pending_nflag_ptr.argtypes = [vp]
    L.bfk_search_pending_nflag_ptr.restype = vp
    L.bfk_xchg_create.argtypes = [i32, i32, i32, i64, ctypes.POINTER(vp), vp]
    L.bfk_xchg_connect.argtypes = [vp, vp]
    L.bfk_xchg_capacity.argtypes = [vp]
    L.bfk_xchg_capacity.restype = i64
    L.bfk_xchg_merge.argtypes = [vp, vp, vp, i64, i64, i32, vp, vp, vp, vp]
    L.bfk_xchg_last_max_nflag.argtypes = [vp]
    L.bfk_xchg_last_max_nflag.restype = i64
    L.bfk_xchg_free.argtypes = [vp]
    L.bfk_xchg_set_id_map.argtypes = [vp, i64]
    L.bfk_reset.argtypes = [vp]
    L.bfk_reconstruct.argtypes = [vp, i64, i64, vp, i32]
    L.bfk_remove_ids.argtypes = [vp, vp, i64, ctypes.POINTER(i64)]
    L.bfk_write.argtypes = [vp, ctypes.c_char_p]
    L.bfk_write_append.argtypes = [vp, ctypes.c_char_p, i64]
    L.bfk_read.argtypes = [ctypes.c_char_p, i32, i32, ctypes.POINTER(vp)]
    L.bfk_serialized_size.argtypes = [vp]
    L.bfk_serialized_size.restype = i64
    L.bfk_serialize.argtypes = [vp, vp, i64]
    L.bfk_deserialize.argtypes = [vp, i64, i32, i32, ctypes.POINTER(vp)]
    L.bfk_normalize_l2.argtypes = [vp, i64, i32, i32, i32, vp]
    L.bfk_merge_topk.argtypes = [vp, vp, i32, i64, i64, i32, vp, vp, i32, vp]
    L.bfk_merge_topk_strided.argtypes = [vp, i64, vp, i64, i32, i64, i64, i32, vp, vp, i32, vp]
    L.bfk_search_noalloc.argtypes = [vp, vp, i64, i64, vp, vp, vp]
    L.bfk_search_sel.argtypes = [vp, vp, i64, i64, vp, vp, i32, vp, i32, i64, vp, i32, ctypes.POINTER(SearchStats)]
    L.bfk_range_search.argtypes = [vp, vp, i64, ctypes.c_float, i32, vp, vp, ctypes.POINTER(vp)]
    L.bfk_range_result_size.argtypes = [vp]
    L.bfk_range_result_size.restype = i64
    L.bfk_range_result_copy.argtypes = [vp, vp, vp, i32]
    L.bfk_range_result_free.argtypes = [vp]
    L.bfk_enc_create.argtypes = [i32, i32, i32, i32, i32, i32, i32, ctypes.c_float, ctypes.POINTER(vp)]
    L.bfk_enc_free.argtypes = [vp]
    L.bfk_enc_set_tensor.argtypes = [vp, ctypes.c_char_p, vp, i64]
    L.bfk_enc_forward.argtypes = [vp, vp, vp, i32, i64, i32, vp, vp]
    L.bfk_enc_encode.argtypes = [vp, vp, vp, i32, i64, i32, vp, vp]
    L.bfk_enc_encode_add.argtypes = [vp, vp, vp, vp, i32, i64, i32, vp]
    _lib = L
    return L


def check(rc: int) -> None:
    """Non-zero status -> RuntimeError, as faiss's FaissException surfaces in Python."""
    if rc != 0:
        msg = load().bfk_last_error()
        raise RuntimeError(msg.decode("utf-8", "replace") if msg else f"libbfk error {rc}")

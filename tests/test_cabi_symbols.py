"""The C-ABI library loads without a GPU and exports every symbol include/bfk.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "bfk.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bfk_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_reference_surface():
    names = declared_symbols()
    for need in ("bfk_create", "bfk_add", "bfk_search", "bfk_ntotal", "bfk_write", "bfk_read",
                 "bfk_pool_normalize", "bfk_add_pooled", "bfk_merge_topk", "bfk_reconstruct", "bfk_remove_ids"):
        assert need in names


def test_library_exports_every_declared_symbol():
    import docqa_b200
    from docqa_b200 import _lib

    L = _lib.load()
    names = declared_symbols()
    assert set(_lib.EXPORTS) <= set(names), "python binding lists a symbol the header does not declare"
    for n in names:
        assert hasattr(L, n), f"libbfk.so does not export {n}"
    assert L.bfk_version() >= 100
    assert docqa_b200.METRIC_L2 == 1 and docqa_b200.METRIC_INNER_PRODUCT == 0


def test_no_cpu_fallback_without_a_gpu():
    from docqa_b200 import _lib
    import docqa_b200

    L = _lib.load()
    if L.bfk_device_count() > 0:
        pytest.skip("a GPU is present")
    h = ctypes.c_void_p()
    assert L.bfk_create(384, 1, 0, 0, ctypes.byref(h)) != 0
    assert b"no CUDA device" in L.bfk_last_error()
    with pytest.raises(RuntimeError, match="no CUDA device"):
        docqa_b200.IndexFlatL2(384)
    import numpy as np

    with pytest.raises(RuntimeError, match="no CUDA device"):
        docqa_b200.pool_normalize(np.zeros((1, 2, 4), np.float32), np.ones((1, 2), np.int64))
    with pytest.raises(RuntimeError, match="no CUDA device"):
        docqa_b200.normalize_L2(np.ones((2, 4), np.float32))
    with pytest.raises(RuntimeError):
        docqa_b200.read_index(os.path.join(ROOT, "tests", "golden", "docqa_vector_store.faiss"))


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "docqa-ms-clinical-document-qa-assistant-llm-microservices-_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "libflat_ref" not in text, f


def test_faiss_shim_exposes_the_names_the_reference_imports():
    from docqa_b200 import faiss_shim

    m = faiss_shim.make_module()
    for name in ("IndexFlatL2", "IndexFlatIP", "IndexFlat", "write_index", "read_index", "normalize_L2",
                 "METRIC_L2", "METRIC_INNER_PRODUCT"):
        assert hasattr(m, name)

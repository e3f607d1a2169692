"""`install_as_faiss()`: make `import faiss` resolve to this package.

semantic-indexer/indexer.py:9, llm-qa/main.py:3 and langchain_community.vectorstores.faiss
import the module by the literal name `faiss`; after install_as_faiss() they get the
B200-backed IndexFlat* / read_index / write_index / normalize_L2 without any edit.
"""
from __future__ import annotations

import sys
import types

from . import flat, idselect


def make_module() -> types.ModuleType:
    m = types.ModuleType("faiss")
    m.__doc__ = "B200-native stand-in for the slice of faiss the DocQA semantic-indexer uses"
    for name in ("IndexFlat", "IndexFlatL2", "IndexFlatIP", "write_index", "read_index", "serialize_index",
                 "deserialize_index", "normalize_L2", "METRIC_L2", "METRIC_INNER_PRODUCT", "get_num_gpus"):
        setattr(m, name, getattr(flat, name))
    for name in ("SearchParameters", "IDSelector", "IDSelectorAll", "IDSelectorRange", "IDSelectorArray",
                 "IDSelectorBatch", "IDSelectorBitmap", "IDSelectorNot", "IDSelectorAnd", "IDSelectorOr"):
        setattr(m, name, getattr(idselect, name))
    m.Index = flat.IndexFlat
    m.__version__ = "bfk-1.0"
    return m


def install_as_faiss(force: bool = False) -> types.ModuleType:
    if "faiss" in sys.modules and not force and getattr(sys.modules["faiss"], "__version__", "") != "bfk-1.0":
        raise RuntimeError("a different `faiss` module is already imported; pass force=True to replace it")
    m = make_module()
    sys.modules["faiss"] = m
    return m

"""B200-native drop-in for the DocQA semantic-indexer's flat vector index.

The directory name follows the build contract and is not a Python identifier; import the
package as `docqa_b200` (a thin alias package at the repository root).
"""
from ._lib import (ALGO_AUTO, ALGO_EXACT_SIMT, ALGO_TENSOR, MAX_D, MAX_K, METRIC_INNER_PRODUCT, METRIC_L2,
                   SearchStats)
from .encoder import MiniLMEncoder, pad_to_multiple
from .faiss_shim import install_as_faiss
from .flat import (IndexFlat, IndexFlatIP, IndexFlatL2, deserialize_index, get_num_gpus, normalize_L2,
                   read_index, serialize_index, write_index, write_index_append)
from .idselect import (IDSelector, IDSelectorAll, IDSelectorAnd, IDSelectorArray, IDSelectorBatch, IDSelectorBitmap,
                       IDSelectorNot, IDSelectorOr, IDSelectorRange, SearchParameters)
from .indexer_host import SemanticIndexer
from .pooling import pool_normalize
from .search_service import create_app, patient_snippets
from .vectorstore_host import Document, FlatVectorStore, InMemoryDocstore
from .sharded import ShardedIndexFlat, shard_range

__all__ = [
    "IndexFlat", "IndexFlatL2", "IndexFlatIP", "write_index", "write_index_append", "read_index", "serialize_index",
    "deserialize_index", "normalize_L2", "METRIC_L2", "METRIC_INNER_PRODUCT", "pool_normalize",
    "install_as_faiss", "SemanticIndexer", "ShardedIndexFlat", "shard_range", "get_num_gpus", "SearchStats",
    "ALGO_AUTO", "ALGO_EXACT_SIMT", "ALGO_TENSOR", "MAX_K", "MAX_D",
    "SearchParameters", "IDSelector", "IDSelectorAll", "IDSelectorRange", "IDSelectorArray", "IDSelectorBatch",
    "IDSelectorBitmap", "IDSelectorNot", "IDSelectorAnd", "IDSelectorOr",
    "MiniLMEncoder", "pad_to_multiple", "FlatVectorStore", "Document", "InMemoryDocstore", "create_app", "patient_snippets",
]

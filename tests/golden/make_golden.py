"""Regenerates tests/golden/* from the reference checkout (run in the CPU container).

    python tests/golden/make_golden.py [/root/reference]

Inputs (reference artefacts of the hot path):
    semantic-indexer/vector_store.faiss   -- written by indexer.py:27
    semantic-indexer/metadata_store.pkl   -- written by indexer.py:28-29
Outputs:
    docqa_vector_store.faiss     byte copy of the index artefact (data, not source):
                                 the known-answer file for write_index/read_index
    docqa_metadata_manifest.json per-row source/type/doc_id + sha1 of the text
                                 (the texts themselves are not copied)
    docqa_search_golden.npz      queries (the 649 corpus rows + 64 seeded perturbed
                                 rows) and float64-exact top-5 L2 answers computed by
                                 oracle/flat_ref.py::knn_f64.  faiss itself cannot run
                                 here, so these pin the oracle, not faiss.
"""
import hashlib
import json
import os
import pickle
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import faiss_io, flat_ref  # noqa: E402


def main(ref="/root/reference"):
    src = os.path.join(ref, "semantic-indexer", "vector_store.faiss")
    raw = open(src, "rb").read()
    with open(os.path.join(HERE, "docqa_vector_store.faiss"), "wb") as f:
        f.write(raw)
    d, n, metric, x = faiss_io.read_index_bytes(raw)
    assert (d, n, metric) == (384, 649, 1)
    assert faiss_io.write_index_bytes(x, metric) == raw

    with open(os.path.join(ref, "semantic-indexer", "metadata_store.pkl"), "rb") as f:
        meta = pickle.load(f)
    assert len(meta) == n
    manifest = {
        "index_sha256": hashlib.sha256(raw).hexdigest(),
        "rows": [
            {
                "doc_id": m["doc_id"],
                "source": m["source"],
                "type": m["type"],
                "text_sha1": hashlib.sha1(m["text_content"].encode("utf-8")).hexdigest(),
                "text_len": len(m["text_content"]),
            }
            for m in meta
        ],
    }
    with open(os.path.join(HERE, "docqa_metadata_manifest.json"), "w") as f:
        json.dump(manifest, f, indent=0)

    rng = np.random.default_rng(4321)
    rows = rng.integers(0, n, size=64)
    pert = x[rows] + 0.05 * rng.standard_normal((64, d)).astype(np.float32)
    pert /= np.linalg.norm(pert, axis=1, keepdims=True)
    q = np.concatenate([x, pert.astype(np.float32)], axis=0)
    D64, I = flat_ref.knn_f64(q, x, 5, flat_ref.METRIC_L2)
    np.savez_compressed(
        os.path.join(HERE, "docqa_search_golden.npz"),
        perturbed_queries=pert.astype(np.float32),
        perturbed_rows=rows,
        D64=D64,
        I=I,
    )
    print("golden written:", len(raw), "bytes index,", len(meta), "rows,", q.shape[0], "queries")


if __name__ == "__main__":
    main(*sys.argv[1:])

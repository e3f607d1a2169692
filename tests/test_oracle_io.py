"""Oracle vs the reference's shipped artefact: file format is byte-exact (not gpu)."""
import hashlib
import json
import os
import struct

import numpy as np

from oracle import faiss_io


def test_fixture_header_fields(fixture_index_bytes):
    b = fixture_index_bytes
    assert len(b) == 996909 == 45 + 649 * 384 * 4
    assert b[:4] == b"IxF2"
    assert struct.unpack_from("<i", b, 4)[0] == 384
    assert struct.unpack_from("<q", b, 8)[0] == 649
    assert struct.unpack_from("<q", b, 16)[0] == 1 << 20
    assert struct.unpack_from("<q", b, 24)[0] == 1 << 20
    assert b[32] == 1
    assert struct.unpack_from("<i", b, 33)[0] == 1
    assert struct.unpack_from("<Q", b, 37)[0] == 649 * 384
    assert b[:45].hex() == ("49784632" "80010000" "8902000000000000" "0000100000000000" "0000100000000000"
                            "01" "01000000" "80cd030000000000")


def test_fixture_round_trip_is_byte_identical(fixture_index_bytes):
    d, n, metric, x = faiss_io.read_index_bytes(fixture_index_bytes)
    assert (d, n, metric) == (384, 649, faiss_io.METRIC_L2)
    assert x.dtype == np.float32 and x.shape == (649, 384)
    assert faiss_io.write_index_bytes(x, metric) == fixture_index_bytes


def test_fixture_matches_manifest(golden_dir, fixture_index_bytes):
    man = json.load(open(os.path.join(golden_dir, "docqa_metadata_manifest.json")))
    assert man["index_sha256"] == hashlib.sha256(fixture_index_bytes).hexdigest()
    rows = man["rows"]
    assert len(rows) == 649  # positional join: index row i <-> metadata_store[i] (indexer.py:41-48)
    assert {r["doc_id"] for r in rows} == {"KB_MTC"}  # indexer.py:44 hard-codes it
    assert [r["type"] for r in rows[:647]] == ["knowledge_base"] * 647
    assert [r["type"] for r in rows[647:]] == ["patient_file"] * 2
    # the same patient document ingested twice -> identical text -> identical vectors
    assert rows[647]["text_sha1"] == rows[648]["text_sha1"]
    _, _, _, x = faiss_io.read_index_bytes(fixture_index_bytes)
    assert (x[647] == x[648]).all()
    # every duplicate text has a duplicate vector
    by_text = {}
    for i, r in enumerate(rows):
        by_text.setdefault(r["text_sha1"], []).append(i)
    for ids in by_text.values():
        for j in ids[1:]:
            assert (x[ids[0]] == x[j]).all()


def test_fixture_rows_are_unit_norm(fixture_corpus):
    nr = np.linalg.norm(fixture_corpus.astype(np.float64), axis=1)
    assert np.abs(nr - 1).max() < 2e-7
    assert len(np.unique(fixture_corpus, axis=0)) == 625  # 22 duplicate pairs + 1 triple


def test_ip_and_legacy_fourcc():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((7, 5)).astype(np.float32)
    b = faiss_io.write_index_bytes(x, faiss_io.METRIC_INNER_PRODUCT)
    assert b[:4] == b"IxFI" and struct.unpack_from("<i", b, 33)[0] == 0
    d, n, metric, y = faiss_io.read_index_bytes(b)
    assert (d, n, metric) == (5, 7, 0) and (x == y).all()
    legacy = b"IxFl" + b[4:]
    d, n, metric, y = faiss_io.read_index_bytes(legacy)
    assert metric == 0 and (x == y).all()
    empty = faiss_io.write_index_bytes(np.zeros((0, 384), np.float32), faiss_io.METRIC_L2)
    assert len(empty) == 45
    assert faiss_io.read_index_bytes(empty)[1] == 0


def test_corrupt_files_raise():
    import pytest

    x = np.zeros((3, 4), np.float32)
    b = faiss_io.write_index_bytes(x, 1)
    with pytest.raises(RuntimeError):
        faiss_io.read_index_bytes(b[:-1])
    with pytest.raises(RuntimeError):
        faiss_io.read_index_bytes(b"IVFx" + b[4:])
    with pytest.raises(RuntimeError):
        faiss_io.read_index_bytes(b[:20])

"""Host mirror of semantic-indexer/indexer.py. CPU only: encoder and index are fakes.

Follows the reference's own tests (semantic-indexer/tests/test_indexer.py:34-59) -- encode is called with
the text list, the index is created once, add is called, metadata is appended, blank text is ignored --
and adds the behaviours the reference does not test: chunking, batching, atomic persistence, and the CSV
templates, which must regenerate the texts behind the shipped index (tests/golden manifest).
"""
import hashlib
import json
import os
import pickle

import numpy as np
import pytest

from oracle import faiss_io, flat_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class FakeModel:
    """encode(list[str]) -> deterministic unit vectors; records calls like the reference's MagicMock."""

    def __init__(self, d=8):
        self.d = d
        self.calls = []

    def encode(self, texts):
        self.calls.append(list(texts))
        out = np.empty((len(texts), self.d), np.float32)
        for i, t in enumerate(texts):
            rng = np.random.default_rng(int(hashlib.sha1(t.encode()).hexdigest()[:8], 16))
            v = rng.standard_normal(self.d)
            out[i] = v / np.linalg.norm(v)
        return out


def ref_write(index, path):
    open(path, "wb").write(faiss_io.write_index_bytes(index.reconstruct_n(0, index.ntotal), index.metric_type))


def ref_read(path):
    d, n, metric, x = faiss_io.read_index_bytes(open(path, "rb").read())
    idx = flat_ref.IndexFlatRef(d, metric)
    idx.add(x)
    return idx


def ref_append(index, path, rows_on_disk):
    """What bfk_write_append does, restated on the oracle index: payload tail past the old end, then the header."""
    raw = open(path, "rb").read()
    d, n, metric, _ = faiss_io.read_index_bytes(raw)
    if (d, metric, n) != (index.d, index.metric_type, rows_on_disk):
        raise RuntimeError("header does not match")
    tail = index.reconstruct_n(rows_on_disk, index.ntotal - rows_on_disk)
    new_header = faiss_io.write_index_bytes(np.zeros((0, d), np.float32), metric)[:45]
    full = faiss_io.write_index_bytes(index.reconstruct_n(0, index.ntotal), metric)
    with open(path, "r+b") as f:
        f.seek(45 + 4 * d * rows_on_disk)
        f.write(np.ascontiguousarray(tail, dtype=np.float32).tobytes())
        f.truncate(45 + 4 * d * index.ntotal)
        f.seek(0)
        f.write(full[:45])
    assert len(new_header) == 45


def make(tmp_path, **kw):
    from docqa_b200.indexer_host import SemanticIndexer

    created = []

    def factory(d):
        created.append(d)
        return flat_ref.IndexFlatRef(d)

    ix = SemanticIndexer(FakeModel(), 8, index_file=str(tmp_path / "vector_store.faiss"),
                         metadata_file=str(tmp_path / "metadata_store.pkl"), index_factory=factory,
                         write_index=ref_write, read_index=ref_read, append_index=ref_append, **kw)
    return ix, created


def test_add_to_index_like_the_reference_test(tmp_path):
    ix, created = make(tmp_path)
    ix.add_to_index("Test sentence", "test_source.txt")
    assert ix.model.calls == [["Test sentence"]]          # model.encode([text])
    assert created == [8]                                  # IndexFlatL2(dimension) once
    assert ix.index.ntotal == 1
    assert ix.metadata_store == [{"doc_id": "KB_MTC", "text_content": "Test sentence",
                                  "source": "test_source.txt", "type": "knowledge_base"}]
    ix.add_to_index("second", "s2", "patient_file")
    assert created == [8] and ix.index.ntotal == 2 and ix.metadata_store[1]["type"] == "patient_file"


def test_add_to_index_empty(tmp_path):
    ix, created = make(tmp_path)
    ix.add_to_index("   ", "source")
    assert ix.metadata_store == [] and ix.index is None and ix.model.calls == []


def test_handle_message_chunks_batches_and_persists_atomically(tmp_path):
    ix, _ = make(tmp_path)
    text = "".join(chr(ord("a") + i % 26) for i in range(1203))  # 3 chunks: 500, 500, 203
    assert ix.handle_message(json.dumps({"doc_id": 7, "original_text_masked": text}).encode()) is True
    assert len(ix.model.calls) == 1 and [len(c) for c in ix.model.calls[0]] == [500, 500, 203]  # ONE encode call
    assert ix.index.ntotal == 3
    assert [m["source"] for m in ix.metadata_store] == ["Dossier Patient 7"] * 3
    assert all(m["doc_id"] == "KB_MTC" and m["type"] == "patient_file" for m in ix.metadata_store)
    # both files exist, no temporaries left, and they agree row for row
    assert sorted(os.listdir(tmp_path)) == ["metadata_store.pkl", "vector_store.faiss"]
    meta = pickle.load(open(tmp_path / "metadata_store.pkl", "rb"))
    d, n, metric, x = faiss_io.read_index_bytes(open(tmp_path / "vector_store.faiss", "rb").read())
    assert n == len(meta) == 3 and metric == 1
    np.testing.assert_array_equal(x, FakeModel().encode([m["text_content"] for m in meta]))
    # malformed message -> nack, state untouched
    assert ix.handle_message(b"not json") is False
    assert ix.handle_message(json.dumps({"doc_id": 1}).encode()) is False  # no text: the reference nacks too
    assert ix.index.ntotal == 3
    # restart: load_or_build picks the files up (indexer.py:97-101)
    ix2, created2 = make(tmp_path)
    ix2.load_or_build()
    assert created2 == [] and ix2.index.ntotal == 3 and ix2.metadata_store == meta


def test_csv_templates_regenerate_the_shipped_corpus(tmp_path):
    ref = "/root/reference/semantic-indexer/default_data"
    if not os.path.isdir(ref):
        pytest.skip("reference checkout not present (GPU box)")
    ix, _ = make(tmp_path, default_data_dir=ref)
    n = ix.ingest_csv_data(batch_rows=64)
    man = json.load(open(os.path.join(ROOT, "tests", "golden", "docqa_metadata_manifest.json")))["rows"]
    assert n == 647 == ix.index.ntotal
    assert max(len(c) for c in ix.model.calls) == 64       # batched encode calls
    for got, want in zip(ix.metadata_store, man[:647]):    # same texts, same order, same sources
        assert hashlib.sha1(got["text_content"].encode("utf-8")).hexdigest() == want["text_sha1"]
        assert got["source"] == want["source"] and got["type"] == want["type"] == "knowledge_base"


def test_device_encoder_path_uses_fused_pooling(tmp_path):
    """An encoder that returns token states is routed to add_pooled (K1 fused with the append)."""
    from docqa_b200.indexer_host import SemanticIndexer

    class TokenModel:
        def encode_tokens(self, texts):
            return ("tok", len(texts)), ("mask", len(texts))

    class PooledIndex:
        ntotal = 0

        def add_pooled(self, tok, mask):
            assert tok[0] == "tok" and mask[0] == "mask"
            self.ntotal += tok[1]

    ix = SemanticIndexer(TokenModel(), 8, index_factory=lambda d: PooledIndex(),
                         index_file=str(tmp_path / "i"), metadata_file=str(tmp_path / "m"))
    assert ix.add_many(["a", " ", "b"], "src") == 2
    assert ix.index.ntotal == 2 and len(ix.metadata_store) == 2


def test_save_state_appends_instead_of_rewriting(tmp_path):
    """indexer.py:125 saves after every message; only the first save writes whole files."""
    ix, _ = make(tmp_path)
    modes, sizes = [], []
    for m in range(6):
        text = f"message {m} " * 90  # 2 chunks
        assert ix.handle_message(json.dumps({"doc_id": m, "original_text_masked": text}).encode())
        modes.append(ix.last_save_mode)
        sizes.append(os.path.getsize(tmp_path / "vector_store.faiss"))
        # after every save the pair on disk is exactly what a full rewrite of the current state gives
        meta = pickle.load(open(tmp_path / "metadata_store.pkl", "rb"))
        assert meta == ix.metadata_store and isinstance(meta, list)
        raw = open(tmp_path / "vector_store.faiss", "rb").read()
        assert raw == faiss_io.write_index_bytes(ix.index.reconstruct_n(0, ix.index.ntotal), 1)
    assert modes == ["full"] + ["append"] * 5
    assert sizes == [45 + 32 * 2 * (m + 1) for m in range(6)]
    assert sorted(os.listdir(tmp_path)) == ["metadata_store.pkl", "vector_store.faiss"]
    ix.save_state()
    assert ix.last_save_mode == "noop"
    # restart: the files are in the appendable layout, so the next message appends straight away
    ix2, _ = make(tmp_path)
    ix2.load_or_build()
    assert ix2.metadata_store == ix.metadata_store and ix2.index.ntotal == 12
    assert ix2.handle_message(json.dumps({"doc_id": 9, "original_text_masked": "x" * 10}).encode())
    assert ix2.last_save_mode == "append"
    assert len(pickle.load(open(tmp_path / "metadata_store.pkl", "rb"))) == 13
    # rows removed: the file no longer holds a prefix of the index -> full rewrite, then appends again
    ix2.index.remove_ids(np.array([0, 5]))
    del ix2.metadata_store[5], ix2.metadata_store[0]
    ix2.invalidate_saved_state()
    ix2.save_state()
    assert ix2.last_save_mode == "full"
    d, n, metric, x = faiss_io.read_index_bytes(open(tmp_path / "vector_store.faiss", "rb").read())
    assert n == 11 == len(pickle.load(open(tmp_path / "metadata_store.pkl", "rb")))
    # checkpoint_every: every third save is a full atomic rewrite
    ix3, _ = make(tmp_path, checkpoint_every=3)
    ix3.load_or_build()
    got = []
    for m in range(6):
        ix3.handle_message(json.dumps({"doc_id": m, "original_text_masked": "y" * 5}).encode())
        got.append(ix3.last_save_mode)
    assert got == ["append", "append", "full", "append", "append", "full"]


def test_metadata_written_by_the_reference_is_rewritten_once_then_appended(tmp_path):
    """A pair of files as the reference leaves them (pickle.dump of the whole list, indexer.py:29-30)."""
    ix, _ = make(tmp_path)
    ix.add_many(["a", "b", "c"], "src")
    ref_write(ix.index, str(tmp_path / "vector_store.faiss"))
    with open(tmp_path / "metadata_store.pkl", "wb") as f:
        pickle.dump(ix.metadata_store, f)
    ix2, _ = make(tmp_path)
    ix2.load_or_build()
    ix2.add_to_index("d", "src")
    ix2.save_state()
    assert ix2.last_save_mode == "full"
    ix2.add_to_index("e", "src")
    ix2.save_state()
    assert ix2.last_save_mode == "append"
    assert [m["text_content"] for m in pickle.load(open(tmp_path / "metadata_store.pkl", "rb"))] == list("abcde")


def test_interrupted_append_leaves_a_loadable_pair(tmp_path):
    """The two commit points: bytes written past the old end are invisible until the one small write that
    publishes them; metadata longer than the index (crash between the two commits) is cut back on load."""
    from docqa_b200 import indexer_host

    ix, _ = make(tmp_path)
    ix.add_many(["a", "b"], "src")
    ix.save_state()
    before = pickle.load(open(tmp_path / "metadata_store.pkl", "rb"))
    # (1) the metadata group has been written but not yet published: readers still see the old list
    with open(tmp_path / "metadata_store.pkl", "ab") as f:
        f.write(indexer_host._group_bytes([{"text_content": "c"}])[1:] + b".")
    assert pickle.load(open(tmp_path / "metadata_store.pkl", "rb")) == before
    # (2) index payload appended, header not yet rewritten: readers still see two rows
    with open(tmp_path / "vector_store.faiss", "ab") as f:
        f.write(np.ones(8, np.float32).tobytes())
    d, n, metric, x = faiss_io.read_index_bytes(open(tmp_path / "vector_store.faiss", "rb").read()[: 45 + 64])
    assert n == 2
    # a process that restarts on such a pair appends behind the real end of the pickle, not behind the debris
    ixr, _ = make(tmp_path)
    ixr._read_index = ref_read_prefix
    ixr.load_or_build()
    assert ixr.index.ntotal == 2 and ixr.metadata_store == before
    ixr.add_to_index("z", "src")
    ixr.save_state()
    assert ixr.last_save_mode == "append"
    assert [m["text_content"] for m in pickle.load(open(tmp_path / "metadata_store.pkl", "rb"))] == ["a", "b", "z"]
    raw = open(tmp_path / "vector_store.faiss", "rb").read()
    assert raw == faiss_io.write_index_bytes(ixr.index.reconstruct_n(0, 3), 1)  # debris gone, byte-identical
    # (3) metadata committed, index not: load cuts the metadata back and schedules a full rewrite
    indexer_host.append_metadata(str(tmp_path / "metadata_store.pkl"), [{"text_content": "orphan"}])
    ix2, _ = make(tmp_path)
    ix2.load_or_build()
    assert ix2.index.ntotal == 3 and len(ix2.metadata_store) == 3
    ix2.add_to_index("d", "src")
    ix2.save_state()
    assert ix2.last_save_mode == "full"
    assert [m["text_content"] for m in pickle.load(open(tmp_path / "metadata_store.pkl", "rb"))] == ["a", "b", "z", "d"]


def ref_read_prefix(path):
    """faiss's reader: header, then exactly `count` floats; bytes after them are not looked at."""
    raw = open(path, "rb").read()
    n = int(np.frombuffer(raw[8:16], dtype=np.int64)[0])
    d = int(np.frombuffer(raw[4:8], dtype=np.int32)[0])
    return ref_read_bytes(raw[: 45 + 4 * n * d])


def ref_read_bytes(raw):
    d, n, metric, x = faiss_io.read_index_bytes(raw)
    idx = flat_ref.IndexFlatRef(d, metric)
    idx.add(x)
    return idx

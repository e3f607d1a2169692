"""Host mirrors either side of the index, CPU only: the oracle index is injected.

  * vectorstore_host.FlatVectorStore -- the LangChain FAISS store llm-qa/main.py:30-59,98-104 builds;
  * search_service -- the patient-snippets endpoint synthese-comparative/core/retrieval_client.py:58-91 calls.
"""
import json
import pickle

import numpy as np
import pytest

from oracle import flat_ref
from tests.test_indexer_host import FakeModel, make, ref_read, ref_write


class Emb:
    """HuggingFaceEmbeddings-shaped wrapper over the fake model."""

    def __init__(self, model):
        self.model = model

    def embed_query(self, text):
        return self.model.encode([text])[0].tolist()

    def embed_documents(self, texts):
        return [v.tolist() for v in self.model.encode(list(texts))]


def build_files(tmp_path, n_kb=30):
    ix, _ = make(tmp_path)
    ix.add_many([f"plante {i} tonifie le qi" for i in range(n_kb)], "herbs.csv")
    ix.handle_message(json.dumps({"doc_id": "D1", "original_text_masked": "patient fatigue " * 80}))
    return ix


def test_store_rebuilds_the_index_text_link_like_llm_qa(tmp_path):
    from docqa_b200.vectorstore_host import Document, FlatVectorStore

    ix = build_files(tmp_path)
    store = FlatVectorStore.from_indexer_files(Emb(ix.model), ix.index_file, ix.metadata_file, read_index=ref_read)
    assert store.index.ntotal == len(ix.metadata_store) == len(store.index_to_docstore_id)
    doc = store.docstore.search("3")
    assert isinstance(doc, Document) and doc.page_content == ix.metadata_store[3]["text_content"]
    assert doc.metadata == {"source": "herbs.csv", "original_id": "KB_MTC", "type": "knowledge_base"}
    # the retriever llm-qa/main.py:101 builds: k=3 nearest documents, best first
    retriever = store.as_retriever(search_kwargs={"k": 3})
    q = "plante 7 tonifie le qi"
    docs = retriever.invoke(q)
    assert len(docs) == 3 and docs[0].page_content == q
    D, I = store.index.search(np.array([Emb(ix.model).embed_query(q)], dtype=np.float32), 3)
    assert [d.page_content for d in docs] == [ix.metadata_store[i]["text_content"] for i in I[0]]
    scored = store.similarity_search_with_score(q, k=3)
    assert [s for _, s in scored] == pytest.approx(D[0].tolist()) and scored[0][1] == pytest.approx(0.0, abs=1e-6)


def test_filter_threshold_and_exact_filter(tmp_path):
    from docqa_b200.vectorstore_host import FlatVectorStore

    ix = build_files(tmp_path)
    store = FlatVectorStore.from_indexer_files(Emb(ix.model), ix.index_file, ix.metadata_file, read_index=ref_read)
    q = "plante 2 tonifie le qi"
    # LangChain semantics: search fetch_k rows, filter afterwards -> may come back short
    post = store.similarity_search(q, k=4, filter={"type": "patient_file"}, fetch_k=5)
    assert all(d.metadata["type"] == "patient_file" for d in post) and len(post) < 4
    # the selector route returns the true top-k of the admitted rows
    exact = store.similarity_search_with_score(q, k=2, filter={"type": ["patient_file"]}, exact_filter=True)
    assert len(exact) == 2 and all(d.metadata["type"] == "patient_file" for d, _ in exact)
    rows = [i for i, m in enumerate(ix.metadata_store) if m["type"] == "patient_file"]
    x = store.index.reconstruct_n(0, store.index.ntotal)
    qv = np.array(Emb(ix.model).embed_query(q), dtype=np.float32)
    d2 = ((x[rows] - qv) ** 2).sum(1)
    assert [s for _, s in exact] == pytest.approx(np.sort(d2)[:2].tolist(), rel=1e-5)
    # callable filter + score threshold (L2: keep scores <= threshold)
    near = store.similarity_search_with_score(q, k=10, filter=lambda m: m["source"] == "herbs.csv", fetch_k=30,
                                              score_threshold=1.0)
    assert near and all(s <= 1.0 for _, s in near)
    # the retriever's threshold is a RELEVANCE in [0, 1] (1 - d / sqrt(2) for L2), kept when >= threshold, as
    # VectorStoreRetriever -> similarity_search_with_relevance_scores does; the raw-distance sense above is the
    # opposite one
    th = store.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.999, "k": 5})
    assert [d.page_content for d in th.invoke(q)] == [q]
    rel = store.similarity_search_with_relevance_scores(q, k=5)
    raw = store.similarity_search_with_score(q, k=5)
    assert [r for _, r in rel] == pytest.approx([1.0 - s / np.sqrt(2.0) for _, s in raw])
    assert rel[0][1] == pytest.approx(1.0, abs=1e-6) and all(a[1] >= b[1] for a, b in zip(rel, rel[1:]))
    loose = store.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.0, "k": 5})
    assert len(loose.invoke(q)) == len([1 for _, r in rel if r >= 0.0])
    with pytest.raises(ValueError):
        store.as_retriever(search_type="bogus")


def test_mmr_add_delete_merge(tmp_path):
    from docqa_b200.vectorstore_host import FlatVectorStore, InMemoryDocstore, maximal_marginal_relevance

    model = FakeModel()
    store = FlatVectorStore(Emb(model), flat_ref.IndexFlatRef(8), InMemoryDocstore({}), {})
    ids = store.add_texts([f"t{i}" for i in range(12)], metadatas=[{"n": i} for i in range(12)])
    assert len(ids) == 12 and store.index.ntotal == 12
    # lambda_mult=1 is plain similarity ranking; smaller values trade relevance for diversity
    emb = Emb(model).embed_query("t3")
    plain = store.max_marginal_relevance_search_by_vector(emb, k=4, fetch_k=12, lambda_mult=1.0)
    assert [d.page_content for d in plain] == [d.page_content for d in store.similarity_search_by_vector(emb, k=4)]
    x = store.index.reconstruct_n(0, 12)
    picked = maximal_marginal_relevance(np.array(emb), list(x), lambda_mult=0.3, k=5)
    assert picked[0] == 3 and len(set(picked)) == 5
    mm = store.max_marginal_relevance_search("t3", k=5, fetch_k=12, lambda_mult=0.3)
    assert mm[0].page_content == "t3" and len(mm) == 5
    # delete shifts later rows down, exactly like index.remove_ids
    store.delete([ids[0], ids[5]])
    assert store.index.ntotal == 10 and store.index_to_docstore_id[0] == ids[1]
    assert store.similarity_search("t6", k=1)[0].page_content == "t6"
    with pytest.raises(ValueError):
        store.delete(["nope"])
    other = FlatVectorStore(Emb(model), flat_ref.IndexFlatRef(8), InMemoryDocstore({}), {})
    other.add_texts(["u0", "u1"])
    store.merge_from(other)
    assert store.index.ntotal == 12 and other.index.ntotal == 0
    assert store.similarity_search("u1", k=1)[0].page_content == "u1"


def patient_indexer(tmp_path):
    ix, _ = make(tmp_path, keep_doc_ids=True)
    ix.add_many([f"plante {i}" for i in range(20)], "herbs.csv")
    msgs = [
        {"doc_id": "DOC-1", "patient_id": "P1", "date": "2024-01-10", "original_text_masked": "inr stable " * 60},
        {"doc_id": "DOC-2", "patient_id": "P1", "date": "2024-03-05", "original_text_masked": "episode hemorragique " * 40},
        {"doc_id": "DOC-3", "patient_id": "P2", "date": "2024-02-01", "original_text_masked": "autre patient " * 50},
    ]
    for m in msgs:
        assert ix.handle_message(json.dumps(m))
    return ix


def test_patient_snippets_handler(tmp_path):
    from docqa_b200.search_service import patient_row_mask, patient_snippets

    ix = patient_indexer(tmp_path)
    meta = pickle.load(open(ix.metadata_file, "rb"))
    assert meta[-1]["doc_id"] == "DOC-3" and meta[-1]["patient_id"] == "P2" and meta[0]["doc_id"] == "KB_MTC"
    assert patient_row_mask(meta, "P1").sum() == 2 + 2  # 660 and 840 characters -> 2 chunks each
    out = patient_snippets(ix, "P1")
    assert [o["doc_id"] for o in out] == ["DOC-1", "DOC-1", "DOC-2", "DOC-2"] and set(out[0]) == {"doc_id", "text"}
    assert [o["doc_id"] for o in patient_snippets(ix, "P1", from_date="2024-02-01")] == ["DOC-2", "DOC-2"]
    assert patient_snippets(ix, "P1", to_date="2023-12-31") == [] and patient_snippets(ix, "P9") == []
    # focus: nearest snippets among the patient's rows only, even though a knowledge-base row is the global nearest
    focus = "plante 4"
    near = patient_snippets(ix, "P2", focus=focus, limit=1)
    assert len(near) == 1 and near[0]["doc_id"] == "DOC-3"
    # rows indexed the reference's way (doc_id "KB_MTC") are still found through their source string
    (tmp_path / "b").mkdir()
    ref_style, _ = make(tmp_path / "b")
    ref_style.handle_message(json.dumps({"doc_id": "P7", "original_text_masked": "x" * 10}))
    got = patient_snippets(ref_style, "P7")
    assert len(got) == 1 and got[0]["doc_id"] == "Dossier Patient P7#0"


def test_patient_snippets_http(tmp_path):
    pytest.importorskip("httpx")
    from fastapi.testclient import TestClient

    from docqa_b200.search_service import create_app

    ix = patient_indexer(tmp_path)
    client = TestClient(create_app(ix))
    assert client.get("/health").json() == {"status": "ok", "service": "semantic-indexer", "vectors": ix.index.ntotal}
    # exactly the call core/retrieval_client.py:85-90 makes (None parameters are dropped by httpx)
    r = client.get("/api/search/patient-snippets", params={"patient_id": "P1", "from_date": "2024-01-01",
                                                           "to_date": "2024-12-31", "focus": "episode hemorragique"})
    assert r.status_code == 200
    body = r.json()
    assert isinstance(body, list) and body and all(set(b) == {"doc_id", "text"} for b in body)
    assert {b["doc_id"] for b in body} == {"DOC-1", "DOC-2"} and body[0]["doc_id"] == "DOC-2"
    assert client.get("/api/search/patient-snippets").status_code == 422  # patient_id is required

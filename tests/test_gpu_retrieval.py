"""The host mirrors either side of the index, on the real GPU index. GPU only."""
import json

import numpy as np
import pytest

from oracle import flat_ref
from tests.test_indexer_host import FakeModel
from tests.test_retrieval_host import Emb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def bfk():
    import docqa_b200

    assert docqa_b200.get_num_gpus() >= 1, "GPU tests need a CUDA device"
    return docqa_b200


def test_vector_store_on_the_reference_corpus(bfk, fixture_corpus):
    """llm-qa's store over the shipped 649x384 corpus: retriever answers equal the oracle's, filters are exact."""
    n = fixture_corpus.shape[0]
    index = bfk.IndexFlatL2(384)
    index.add(fixture_corpus)
    docstore = bfk.InMemoryDocstore({str(i): bfk.Document(f"chunk {i}", {"source": "a.csv" if i % 3 else "b.csv", "row": i})
                                     for i in range(n)})
    rows = {f"chunk {i}": fixture_corpus[i] for i in range(n)}
    store = bfk.FlatVectorStore(lambda text: rows[text].tolist(), index, docstore, {i: str(i) for i in range(n)})
    ref = flat_ref.IndexFlatRef(384)
    ref.add(fixture_corpus)
    for r in (0, 74, 321, 648):
        Dr, Ir = ref.search(fixture_corpus[r : r + 1], 3)
        got = store.similarity_search_with_score(f"chunk {r}", k=3)
        assert [s for _, s in got] == pytest.approx(Dr[0].tolist(), abs=1e-6)
        dr64 = flat_ref.pairwise_f64(fixture_corpus[r : r + 1], fixture_corpus, flat_ref.METRIC_L2)[0]
        for (doc, s), want in zip(got, Ir[0]):  # ids equal outside exact ties (the corpus has duplicates)
            assert doc.metadata["row"] == want or abs(dr64[doc.metadata["row"]] - dr64[want]) < 1e-6
        # exact filter: true top-3 of the admitted third of the corpus
        ex = store.similarity_search_with_score(f"chunk {r}", k=3, filter={"source": "b.csv"}, exact_filter=True)
        sub = np.arange(0, n, 3)
        want = np.sort(dr64[sub])[:3]
        assert [s for _, s in ex] == pytest.approx(want.tolist(), abs=1e-5)
        assert all(d.metadata["source"] == "b.csv" for d, _ in ex)
    mm = store.max_marginal_relevance_search("chunk 5", k=4, fetch_k=16, lambda_mult=0.5)
    assert len(mm) == 4 and mm[0].metadata["row"] in (5, 13)  # rows 5 and 13 are duplicates in the fixture


def test_patient_snippets_on_gpu_index(bfk, tmp_path):
    model = FakeModel(d=64)
    ix = bfk.SemanticIndexer(model, 64, index_file=str(tmp_path / "vector_store.faiss"),
                             metadata_file=str(tmp_path / "metadata_store.pkl"), keep_doc_ids=True)
    ix.add_many([f"plante {i}" for i in range(500)], "herbs.csv")
    for p in range(6):
        ix.handle_message(json.dumps({"doc_id": f"DOC-{p}", "patient_id": f"P{p % 3}", "date": f"2024-0{p + 1}-01",
                                      "original_text_masked": f"note {p} " * 200}))
    assert isinstance(ix.index, bfk.IndexFlatL2) and ix.index.ntotal == len(ix.metadata_store)
    got = bfk.patient_snippets(ix, "P1", focus="plante 3", limit=3)
    assert len(got) == 3 and {g["doc_id"] for g in got} <= {"DOC-1", "DOC-4"}
    # same answer as a brute-force pass over that patient's rows
    rows = [i for i, m in enumerate(ix.metadata_store) if m.get("patient_id") == "P1"]
    x = ix.index.reconstruct_n(0, ix.index.ntotal)
    q = model.encode(["plante 3"])[0]
    order = np.argsort(((x[rows] - q) ** 2).sum(1), kind="stable")[:3]
    assert [g["text"] for g in got] == [ix.metadata_store[rows[i]]["text_content"] for i in order]
    assert [g["doc_id"] for g in bfk.patient_snippets(ix, "P1", from_date="2024-03-01")][:1] == ["DOC-4"]
    # files round-trip into llm-qa's store
    store = bfk.FlatVectorStore.from_indexer_files(Emb(model), ix.index_file, ix.metadata_file)
    assert store.index.ntotal == ix.index.ntotal
    assert store.similarity_search("plante 42", k=1)[0].page_content == "plante 42"

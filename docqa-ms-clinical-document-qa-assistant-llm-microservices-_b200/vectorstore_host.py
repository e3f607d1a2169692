"""Host mirror of the LangChain vector store llm-qa builds around the flat index.

llm-qa/main.py:30-59 reads `vector_store.faiss` + `metadata_store.pkl`, wraps row i as
`Document(page_content=meta["text_content"], metadata={"source": ..., "original_id": ...})` under docstore
id str(i), and builds `langchain_community.vectorstores.FAISS(embedding_function, index, docstore,
index_to_docstore_id)`; llm-qa/main.py:98-104 then retrieves through
`vector_store.as_retriever(search_kwargs={"k": 3})`.  langchain_community is not installable here, so this
module restates the slice of that class the path can reach -- same method names, argument meaning and
return shapes -- on top of any index with the faiss.IndexFlat surface (this package's on a B200; the
CPU oracle in the tests):

    similarity_search_with_score_by_vector / similarity_search_with_score / similarity_search
    max_marginal_relevance_search_with_score_by_vector / max_marginal_relevance_search (+ _by_vector)
    as_retriever(search_type=..., search_kwargs=...).invoke(query) / get_relevant_documents(query)
    add_texts / add_embeddings / delete / merge_from / from_indexer_files

Restated LangChain behaviour: scores are the index's raw distances (squared L2 ascending, or inner products
descending for MAX_INNER_PRODUCT); `filter` is a metadata dict (value or list of admitted values per key) or a
callable on the metadata; with a filter LangChain searches `fetch_k` rows and filters afterwards, so it may
return fewer than k -- kept as the default; `score_threshold` keeps scores <= threshold (>= for inner
product); -1 ids are skipped.  Extension: `exact_filter=True` lowers the filter to an id selector
(bfk_search_sel), which returns the true top-k of the admitted rows in one pass.
"""
from __future__ import annotations

import math
import operator
import pickle
import uuid
from dataclasses import dataclass, field

import numpy as np

from .idselect import IDSelectorBitmap, SearchParameters

EUCLIDEAN_DISTANCE = "EUCLIDEAN_DISTANCE"
MAX_INNER_PRODUCT = "MAX_INNER_PRODUCT"
COSINE = "COSINE"


@dataclass
class Document:
    page_content: str
    metadata: dict = field(default_factory=dict)


class InMemoryDocstore:
    """langchain_community.docstore.in_memory.InMemoryDocstore: id -> Document."""

    def __init__(self, _dict=None):
        self._dict = dict(_dict or {})

    def add(self, texts: dict) -> None:
        overlapping = set(texts).intersection(self._dict)
        if overlapping:
            raise ValueError(f"Tried to add ids that already exist: {overlapping}")
        self._dict.update(texts)

    def delete(self, ids) -> None:
        missing = set(ids).difference(self._dict)
        if missing:
            raise ValueError(f"Tried to delete ids that does not  exist: {missing}")
        for _id in ids:
            self._dict.pop(_id)

    def search(self, search: str):
        if search not in self._dict:
            return f"ID {search} not found."
        return self._dict[search]


def _filter_fn(filter):
    """LangChain's FAISS._create_filter_func for the plain forms: callable, or {key: value | [values]}."""
    if filter is None:
        return None
    if callable(filter):
        return filter
    if not isinstance(filter, dict):
        raise ValueError(f"filter must be a dict of metadata or a callable, not {type(filter)}")

    def fn(metadata: dict) -> bool:
        for key, want in filter.items():
            have = metadata.get(key)
            if isinstance(want, (list, tuple, set)):
                if have not in want:
                    return False
            elif have != want:
                return False
        return True

    return fn


def cosine_similarity(X, Y) -> np.ndarray:
    X, Y = np.asarray(X, dtype=np.float64), np.asarray(Y, dtype=np.float64)
    if X.size == 0 or Y.size == 0:
        return np.zeros((X.shape[0], Y.shape[0]))
    xn, yn = np.linalg.norm(X, axis=1), np.linalg.norm(Y, axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        sim = (X @ Y.T) / np.outer(xn, yn)
    sim[~np.isfinite(sim)] = 0.0
    return sim


def maximal_marginal_relevance(query_embedding, embedding_list, lambda_mult: float = 0.5, k: int = 4) -> list[int]:
    """langchain_community.vectorstores.utils.maximal_marginal_relevance."""
    if min(k, len(embedding_list)) <= 0:
        return []
    query_embedding = np.asarray(query_embedding)
    if query_embedding.ndim == 1:
        query_embedding = query_embedding[None, :]
    sim_q = cosine_similarity(query_embedding, embedding_list)[0]
    most = int(np.argmax(sim_q))
    idxs, selected = [most], np.array([embedding_list[most]])
    while len(idxs) < min(k, len(embedding_list)):
        best_score, best_i = -np.inf, -1
        sim_sel = cosine_similarity(embedding_list, selected)
        for i, qs in enumerate(sim_q):
            if i in idxs:
                continue
            score = lambda_mult * qs - (1 - lambda_mult) * sim_sel[i].max()
            if score > best_score:
                best_score, best_i = score, i
        idxs.append(best_i)
        selected = np.append(selected, [embedding_list[best_i]], axis=0)
    return idxs


class FlatVectorStore:
    """The FAISS vector store of llm-qa/main.py:53-58 over a flat index."""

    def __init__(self, embedding_function, index, docstore: InMemoryDocstore, index_to_docstore_id: dict,
                 normalize_L2: bool = False, distance_strategy: str = EUCLIDEAN_DISTANCE):
        self.embedding_function = embedding_function
        self.index = index
        self.docstore = docstore
        self.index_to_docstore_id = index_to_docstore_id
        self._normalize_L2 = normalize_L2
        self.distance_strategy = distance_strategy

    # ---- llm-qa/main.py:30-59 ----------------------------------------------------------
    @classmethod
    def from_indexer_files(cls, embedding_function, faiss_path: str, meta_path: str, *, read_index=None, **kw):
        """Rebuild the index <-> text link exactly as llm-qa does at start-up."""
        if read_index is None:
            from .flat import read_index
        raw_index = read_index(faiss_path)
        with open(meta_path, "rb") as f:
            metadata_store = pickle.load(f)
        docstore, index_to_docstore_id = InMemoryDocstore({}), {}
        for i, meta in enumerate(metadata_store):
            doc_id = str(i)
            extra = {k: v for k, v in meta.items() if k not in ("text_content", "source", "doc_id")}
            doc = Document(page_content=meta["text_content"],
                           metadata={"source": meta["source"], "original_id": meta["doc_id"], **extra})
            docstore.add({doc_id: doc})
            index_to_docstore_id[i] = doc_id
        return cls(embedding_function, raw_index, docstore, index_to_docstore_id, **kw)

    # ---- embedding helpers ---------------------------------------------------------------
    def _embed_query(self, text: str):
        f = self.embedding_function
        return f.embed_query(text) if hasattr(f, "embed_query") else f(text)

    def _embed_documents(self, texts):
        f = self.embedding_function
        return f.embed_documents(texts) if hasattr(f, "embed_documents") else [f(t) for t in texts]

    def _vector(self, embedding) -> np.ndarray:
        v = np.array([embedding], dtype=np.float32)
        if self._normalize_L2:
            n = np.linalg.norm(v, axis=1, keepdims=True)
            v = np.where(n > 0, v / np.where(n > 0, n, 1), v).astype(np.float32)
        return v

    def _selector(self, fn):
        """filter -> id selector over the current rows (one host pass over the metadata)."""
        mask = np.zeros(self.index.ntotal, dtype=bool)
        for i, _id in self.index_to_docstore_id.items():
            doc = self.docstore.search(_id)
            if isinstance(doc, Document) and fn(doc.metadata):
                mask[i] = True
        return IDSelectorBitmap.from_mask(mask)

    # ---- similarity search ---------------------------------------------------------------
    def similarity_search_with_score_by_vector(self, embedding, k: int = 4, filter=None, fetch_k: int = 20,
                                               exact_filter: bool = False, **kwargs):
        vector = self._vector(embedding)
        fn = _filter_fn(filter)
        if fn is not None and exact_filter:
            scores, indices = self.index.search(vector, k, params=SearchParameters(sel=self._selector(fn)))
        else:
            scores, indices = self.index.search(vector, k if fn is None else fetch_k)
        docs = []
        for j, i in enumerate(indices[0]):
            if i == -1:
                continue
            _id = self.index_to_docstore_id[int(i)]
            doc = self.docstore.search(_id)
            if not isinstance(doc, Document):
                raise ValueError(f"Could not find document for id {_id}, got {doc}")
            if fn is None or fn(doc.metadata):
                docs.append((doc, float(scores[0][j])))
        score_threshold = kwargs.get("score_threshold")
        if score_threshold is not None:
            cmp = operator.ge if self.distance_strategy == MAX_INNER_PRODUCT else operator.le
            docs = [(doc, s) for doc, s in docs if cmp(s, score_threshold)]
        return docs[:k]

    def similarity_search_with_score(self, query: str, k: int = 4, filter=None, fetch_k: int = 20, **kwargs):
        return self.similarity_search_with_score_by_vector(self._embed_query(query), k, filter=filter,
                                                           fetch_k=fetch_k, **kwargs)

    def similarity_search_by_vector(self, embedding, k: int = 4, filter=None, fetch_k: int = 20, **kwargs):
        return [d for d, _ in self.similarity_search_with_score_by_vector(embedding, k, filter=filter,
                                                                         fetch_k=fetch_k, **kwargs)]

    def similarity_search(self, query: str, k: int = 4, filter=None, fetch_k: int = 20, **kwargs):
        return [d for d, _ in self.similarity_search_with_score(query, k, filter=filter, fetch_k=fetch_k, **kwargs)]

    # ---- maximal marginal relevance --------------------------------------------------------
    def max_marginal_relevance_search_with_score_by_vector(self, embedding, *, k: int = 4, fetch_k: int = 20,
                                                           lambda_mult: float = 0.5, filter=None):
        fn = _filter_fn(filter)
        scores, indices = self.index.search(np.array([embedding], dtype=np.float32),
                                            fetch_k if fn is None else fetch_k * 2)
        if fn is not None:
            kept = []
            for i in indices[0]:
                if i == -1:
                    continue
                doc = self.docstore.search(self.index_to_docstore_id[int(i)])
                if isinstance(doc, Document) and fn(doc.metadata):
                    kept.append(int(i))
            indices = np.array([kept])
        embeddings = [self.index.reconstruct(int(i)) for i in indices[0] if i != -1]
        mmr = maximal_marginal_relevance(np.array([embedding], dtype=np.float32), embeddings, k=k,
                                         lambda_mult=lambda_mult)
        out = []
        for i in mmr:
            row = int(indices[0][i])
            if row == -1:
                continue
            doc = self.docstore.search(self.index_to_docstore_id[row])
            if not isinstance(doc, Document):
                raise ValueError(f"Could not find document for id {self.index_to_docstore_id[row]}, got {doc}")
            pos = int(np.flatnonzero(np.asarray(indices[0]) == row)[0]) if fn is None else None
            out.append((doc, float(scores[0][pos]) if pos is not None else float("nan")))
        return out

    def max_marginal_relevance_search_by_vector(self, embedding, k: int = 4, fetch_k: int = 20,
                                                lambda_mult: float = 0.5, filter=None, **kwargs):
        return [d for d, _ in self.max_marginal_relevance_search_with_score_by_vector(
            embedding, k=k, fetch_k=fetch_k, lambda_mult=lambda_mult, filter=filter)]

    def max_marginal_relevance_search(self, query: str, k: int = 4, fetch_k: int = 20, lambda_mult: float = 0.5,
                                      filter=None, **kwargs):
        return self.max_marginal_relevance_search_by_vector(self._embed_query(query), k=k, fetch_k=fetch_k,
                                                            lambda_mult=lambda_mult, filter=filter)

    # ---- mutation ---------------------------------------------------------------------------
    def add_embeddings(self, text_embeddings, metadatas=None, ids=None):
        texts, embeddings = zip(*text_embeddings) if text_embeddings else ((), ())
        return self._add(list(texts), list(embeddings), metadatas, ids)

    def add_texts(self, texts, metadatas=None, ids=None):
        texts = list(texts)
        return self._add(texts, self._embed_documents(texts), metadatas, ids)

    def _add(self, texts, embeddings, metadatas, ids):
        if not texts:
            return []
        metadatas = metadatas or [{} for _ in texts]
        ids = list(ids) if ids is not None else [str(uuid.uuid4()) for _ in texts]
        if len(set(ids)) != len(ids):
            raise ValueError("Duplicate ids found in the ids list.")
        vectors = np.array(embeddings, dtype=np.float32).reshape(len(texts), -1)
        if self._normalize_L2:
            n = np.linalg.norm(vectors, axis=1, keepdims=True)
            vectors = (vectors / np.where(n > 0, n, 1)).astype(np.float32)
        self.docstore.add({i: Document(page_content=t, metadata=m) for i, t, m in zip(ids, texts, metadatas)})
        start = len(self.index_to_docstore_id)
        self.index.add(vectors)
        self.index_to_docstore_id.update({start + j: _id for j, _id in enumerate(ids)})
        return ids

    def delete(self, ids=None, **kwargs):
        """Remove documents by docstore id: index.remove_ids shifts later rows down, the map is rebuilt."""
        if ids is None:
            raise ValueError("No ids provided to delete.")
        missing = set(ids).difference(self.index_to_docstore_id.values())
        if missing:
            raise ValueError(f"Some specified ids do not exist in the current store. Ids not found: {missing}")
        reversed_index = {_id: idx for idx, _id in self.index_to_docstore_id.items()}
        index_to_delete = {reversed_index[_id] for _id in ids}
        self.index.remove_ids(np.fromiter(index_to_delete, dtype=np.int64))
        self.docstore.delete(ids)
        remaining = [_id for i, _id in sorted(self.index_to_docstore_id.items()) if i not in index_to_delete]
        self.index_to_docstore_id = {i: _id for i, _id in enumerate(remaining)}
        return True

    # ---- relevance scores (what VectorStoreRetriever uses for search_type="similarity_score_threshold") ------
    def _relevance_fn(self):
        """langchain_community FAISS._select_relevance_score_fn: distance -> relevance in [0, 1]."""
        if self.distance_strategy == MAX_INNER_PRODUCT:
            return lambda s: 1.0 - s if s > 0 else -1.0 * s   # _max_inner_product_relevance_score_fn
        if self.distance_strategy == EUCLIDEAN_DISTANCE:
            return lambda s: 1.0 - s / math.sqrt(2)            # _euclidean_relevance_score_fn (unit-norm embeddings)
        if self.distance_strategy == COSINE:
            return lambda s: 1.0 - s
        raise ValueError(f"Unknown distance strategy {self.distance_strategy}")

    def similarity_search_with_relevance_scores(self, query: str, k: int = 4, filter=None, fetch_k: int = 20, **kwargs):
        """(doc, relevance) pairs, relevance in [0, 1]; `score_threshold` keeps relevance >= threshold -- the
        opposite sense of the raw-distance threshold of similarity_search_with_score, as in LangChain."""
        score_threshold = kwargs.pop("score_threshold", None)
        fn = self._relevance_fn()
        docs = [(doc, fn(s)) for doc, s in self.similarity_search_with_score(query, k, filter=filter, fetch_k=fetch_k, **kwargs)]
        if score_threshold is not None:
            docs = [(doc, r) for doc, r in docs if r >= score_threshold]
        return docs

    def merge_from(self, target: "FlatVectorStore") -> None:
        starting_len = len(self.index_to_docstore_id)
        full_info = []
        for i, target_id in target.index_to_docstore_id.items():
            doc = target.docstore.search(target_id)
            if not isinstance(doc, Document):
                raise ValueError("Document should be returned")
            full_info.append((starting_len + i, target_id, doc))
        self.index.merge_from(target.index)
        self.docstore.add({_id: doc for _, _id, doc in full_info})
        self.index_to_docstore_id.update({index: _id for index, _id, _ in full_info})

    # ---- retriever (llm-qa/main.py:101) --------------------------------------------------------
    def as_retriever(self, search_type: str = "similarity", search_kwargs: dict | None = None):
        return Retriever(self, search_type, dict(search_kwargs or {}))


class Retriever:
    """VectorStoreRetriever: `.invoke(query)` is what RetrievalQA calls for its context documents."""

    allowed = ("similarity", "similarity_score_threshold", "mmr")

    def __init__(self, vectorstore: FlatVectorStore, search_type: str, search_kwargs: dict):
        if search_type not in self.allowed:
            raise ValueError(f"search_type of {search_type} not allowed. Valid values are: {self.allowed}")
        if search_type == "similarity_score_threshold" and not isinstance(search_kwargs.get("score_threshold"), float):
            raise ValueError("`score_threshold` is not specified with a float value(0~1) in `search_kwargs`.")
        self.vectorstore, self.search_type, self.search_kwargs = vectorstore, search_type, search_kwargs

    def invoke(self, query: str, **_):
        if self.search_type == "mmr":
            return self.vectorstore.max_marginal_relevance_search(query, **self.search_kwargs)
        if self.search_type == "similarity_score_threshold":  # relevance >= threshold, as VectorStoreRetriever does
            return [d for d, _ in self.vectorstore.similarity_search_with_relevance_scores(query, **self.search_kwargs)]
        return self.vectorstore.similarity_search(query, **self.search_kwargs)

    get_relevant_documents = invoke

"""HTTP search endpoint of the semantic indexer.

synthese-comparative/core/retrieval_client.py:58-91 expects, on the indexer (core/config.py:10-13,
http://semantic-indexer:8003):

    GET /api/search/patient-snippets?patient_id=P123&from_date=...&to_date=...&focus=...
    -> [{"doc_id": "DOC-123", "text": "..."}, ...]

and nothing in the reference implements it (its real mode cannot work today).  This module does, on top of
`indexer_host.SemanticIndexer` (index + metadata_store, positional join as in indexer.py:43-48):

  * rows of the patient: metadata type "patient_file" whose patient_id (kept when the indexer runs with
    keep_doc_ids=True) or doc_id equals `patient_id`, or whose source is "Dossier Patient <patient_id>"
    (what indexer.py:121 writes); `from_date` / `to_date` bound the row's "date" when it has one (ISO strings
    compare lexicographically);
  * with `focus`: the focus text is embedded and the `limit` nearest snippets AMONG THOSE ROWS are returned --
    one selector-restricted search on the GPU (bfk_search_sel), not a post-filter;
  * without `focus`: the patient's snippets in ingestion order.

`patient_snippets()` is the handler as a plain function; `create_app()` wraps it in FastAPI with the same
/health shape the reference services use (llm-qa/main.py:126-128).
"""
from __future__ import annotations

import numpy as np

from .idselect import IDSelectorBitmap, SearchParameters


def patient_row_mask(metadata_store, patient_id: str, from_date: str | None = None,
                     to_date: str | None = None) -> np.ndarray:
    mask = np.zeros(len(metadata_store), dtype=bool)
    source = f"Dossier Patient {patient_id}"
    for i, meta in enumerate(metadata_store):
        if meta.get("type") != "patient_file":
            continue
        if not (meta.get("patient_id") == patient_id or meta.get("doc_id") == patient_id or meta.get("source") == source):
            continue
        date = meta.get("date")
        if date is not None:
            if from_date and str(date) < from_date:
                continue
            if to_date and str(date) > to_date:
                continue
        mask[i] = True
    return mask


def _snippet(meta: dict, row: int) -> dict:
    doc_id = meta.get("doc_id")
    if not doc_id or doc_id == "KB_MTC":
        doc_id = f"{meta.get('source', 'row')}#{row}"
    return {"doc_id": str(doc_id), "text": meta["text_content"]}


def patient_snippets(indexer, patient_id: str, from_date: str | None = None, to_date: str | None = None,
                     focus: str | None = None, limit: int = 10) -> list[dict]:
    """The endpoint's body: list of {"doc_id", "text"} for one patient."""
    metadata_store = indexer.metadata_store
    index = indexer.index
    if index is None or not metadata_store:
        return []
    n = min(index.ntotal, len(metadata_store))  # a snapshot where the files are momentarily out of step
    mask = patient_row_mask(metadata_store[:n], patient_id, from_date, to_date)
    rows = np.flatnonzero(mask)
    if rows.size == 0:
        return []
    if not focus or not focus.strip():
        return [_snippet(metadata_store[int(r)], int(r)) for r in rows[:limit]]
    emb = np.asarray(indexer.model.encode([focus]), dtype=np.float32).reshape(1, -1)
    full = np.zeros(index.ntotal, dtype=bool)
    full[:n] = mask
    k = int(min(limit, rows.size))
    _, I = index.search(emb, k, params=SearchParameters(sel=IDSelectorBitmap.from_mask(full)))
    I = I.cpu().numpy() if hasattr(I, "cpu") else np.asarray(I)
    return [_snippet(metadata_store[int(r)], int(r)) for r in I[0] if r >= 0]


def create_app(indexer, *, default_limit: int = 10):
    from fastapi import FastAPI, Query

    app = FastAPI(title="Semantic Indexer Search")

    @app.get("/api/search/patient-snippets")
    def search_patient_snippets(patient_id: str, from_date: str | None = None, to_date: str | None = None,
                                focus: str | None = None, limit: int = Query(default_limit, ge=1, le=2048)):
        return patient_snippets(indexer, patient_id, from_date, to_date, focus, limit)

    @app.get("/health")
    def health():
        return {"status": "ok", "service": "semantic-indexer",
                "vectors": int(indexer.index.ntotal) if indexer.index is not None else 0}

    return app

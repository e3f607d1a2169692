"""Host-side mirror of `semantic-indexer/indexer.py` on top of the B200 flat index.

SURVEY.md section 8f, row 1 ("batched, atomic ingestion"): same function names, argument
meaning and side effects as the reference module, so its call sites and its unit tests
read the same, with three changes that the reference's structure invites:

  * `add_many` encodes all chunks of a message in ONE `encode` call and appends them with
    ONE `index.add(n)` (reference: one `encode([text])` + one `add(1)` per chunk,
    indexer.py:37-41,120-123);
  * when the encoder hands back token states on the GPU (`encode_tokens`), pooling +
    normalisation + append run fused on the device (`IndexFlat.add_pooled`, kernel K1) and
    the vectors never visit host memory;
  * `save_state` writes both files to temporaries and renames them into place
    (reference: two in-place writes, indexer.py:26-30, so llm-qa can read a torn pair).

What is deliberately kept: blank texts are skipped (indexer.py:35), `doc_id` is the
hard-coded "KB_MTC" (indexer.py:44), ids are positional (row i of the index <->
metadata_store[i]), patient text is cut into 500-character chunks without overlap
(indexer.py:120), the CSV sentence templates (indexer.py:70-74,81-87), and the message
protocol of `callback` (indexer.py:112-129): returns True to ack, False to nack.

RabbitMQ itself (pika) stays outside: `handle_message(body)` is what the reference's
`callback` would call.  The encoder and the index factory are injected, so the logic is
testable without a GPU.
"""
from __future__ import annotations

import csv
import glob
import json
import os
import pickle
import tempfile

import numpy as np

INDEX_FILE = "vector_store.faiss"
METADATA_FILE = "metadata_store.pkl"
DEFAULT_DATA_DIR = "default_data"
CHUNK_CHARS = 500


def _default_index_factory(dimension: int):
    from .flat import IndexFlatL2

    return IndexFlatL2(dimension)


def _default_io():
    from . import flat

    return flat.write_index, flat.read_index


class SemanticIndexer:
    """State and behaviour of the reference's indexer module, as an object."""

    def __init__(self, model, dimension: int = 384, *, index_file: str = INDEX_FILE,
                 metadata_file: str = METADATA_FILE, default_data_dir: str = DEFAULT_DATA_DIR,
                 index_factory=None, write_index=None, read_index=None, keep_doc_ids: bool = False):
        self.model = model                      # .encode(list[str]) -> [n, d]; optional .encode_tokens
        self.dimension = int(dimension)
        self.index_file = index_file
        self.metadata_file = metadata_file
        self.default_data_dir = default_data_dir
        self._index_factory = index_factory or _default_index_factory
        self._write_index, self._read_index = write_index, read_index  # None -> this package's, resolved lazily
        self.index = None
        self.metadata_store: list[dict] = []
        # False: every row carries doc_id "KB_MTC" like the reference (indexer.py:44).  True: patient messages
        # keep their own doc_id plus patient_id / date when the message has them -- what the patient-snippets
        # endpoint (search_service.py; synthese-comparative/core/retrieval_client.py:58-91) filters on.
        self.keep_doc_ids = bool(keep_doc_ids)

    # ---- indexer.py:26-30 ------------------------------------------------------------
    def save_state(self) -> None:
        """Persist index + metadata; each file appears atomically (temp file + rename)."""
        d = os.path.dirname(os.path.abspath(self.index_file)) or "."
        fd, tmp_idx = tempfile.mkstemp(prefix=".vector_store.", suffix=".tmp", dir=d)
        os.close(fd)
        fd, tmp_meta = tempfile.mkstemp(prefix=".metadata_store.", suffix=".tmp", dir=d)
        try:
            (self._write_index or _default_io()[0])(self.index, tmp_idx)
            with os.fdopen(fd, "wb") as f:
                pickle.dump(self.metadata_store, f, protocol=4)
            # metadata first: a reader that sees the new index always finds metadata at least as long
            os.replace(tmp_meta, self.metadata_file)
            os.replace(tmp_idx, self.index_file)
        finally:
            for p in (tmp_idx, tmp_meta):
                if os.path.exists(p):
                    os.unlink(p)

    # ---- indexer.py:32-48 ------------------------------------------------------------
    def add_to_index(self, text, source_name, doc_type="knowledge_base") -> None:
        """One text -> one vector -> one metadata row (the reference's unit of work)."""
        if not text.strip():
            return
        self.add_many([text], source_name, doc_type)

    def add_many(self, texts, source_name, doc_type="knowledge_base", *, doc_id="KB_MTC", extra=None) -> int:
        """All non-blank texts in one encode call and one add; returns how many were added."""
        texts = [t for t in texts if t.strip()]
        if not texts:
            return 0
        if self.index is None:
            self.index = self._index_factory(self.dimension)
        if hasattr(self.model, "encode_tokens") and hasattr(self.index, "add_pooled"):
            token_embeddings, attention_mask = self.model.encode_tokens(texts)
            self.index.add_pooled(token_embeddings, attention_mask)
        else:
            embeddings = self.model.encode(texts)
            self.index.add(np.array(embeddings).astype("float32").reshape(len(texts), -1))
        for text in texts:
            self.metadata_store.append({
                "doc_id": doc_id,
                "text_content": text,
                "source": source_name,
                "type": doc_type,
                **(extra or {}),
            })
        return len(texts)

    # ---- indexer.py:50-94 ------------------------------------------------------------
    @staticmethod
    def csv_row_text(filename: str, row: dict):
        """The sentence the reference builds for one CSV row, or None if the file is not one it knows."""
        if "matrice" in filename or "ranking" in filename:
            return (
                f"ANALYSE SCORE MTC : Syndrome '{row.get('nom_syndrome', '')}'. "
                f"Plante recommandée : {row.get('nom_latin', '')} ({row.get('nom_chinois', '')}). "
                f"Score de pertinence : {row.get('score_role', '0')}."
            )
        if "base" in filename or "connaissance" in filename:
            return (
                f"DÉTAIL CLINIQUE : Syndrome '{row.get('nom_syndrome', '')}'. "
                f"Formule '{row.get('nom_formule', '')}'. "
                f"Plante : {row.get('nom_latin', '')}. "
                f"Rôle : {row.get('role_formule', 'Inconnu')} (Score {row.get('score_role', '')}). "
                f"Description : {row.get('description', '')}"
            )
        return None

    def ingest_csv_data(self, batch_rows: int = 256) -> int:
        """Index every default_data/*.csv; files in sorted order (what the shipped index shows)."""
        if not os.path.exists(self.default_data_dir):
            return 0
        total = 0
        for file_path in sorted(glob.glob(os.path.join(self.default_data_dir, "*.csv"))):
            filename = os.path.basename(file_path)
            try:
                with open(file_path, "r", encoding="utf-8") as f:
                    batch = []
                    for row in csv.DictReader(f):
                        text = self.csv_row_text(filename, row)
                        if text is None:
                            continue
                        batch.append(text)
                        if len(batch) >= batch_rows:
                            total += self.add_many(batch, filename)
                            batch = []
                    total += self.add_many(batch, filename)
            except Exception as e:  # the reference logs and carries on (indexer.py:93-94)
                print(f"Erreur lecture CSV {filename}: {e}")
        return total

    # ---- indexer.py:97-107 -----------------------------------------------------------
    def load_or_build(self) -> None:
        if os.path.exists(self.index_file) and os.path.exists(self.metadata_file):
            self.index = (self._read_index or _default_io()[1])(self.index_file)
            with open(self.metadata_file, "rb") as f:
                self.metadata_store = pickle.load(f)
        else:
            self.index = self._index_factory(self.dimension)
            self.metadata_store = []
            self.ingest_csv_data()
            self.save_state()

    # ---- indexer.py:112-129 ----------------------------------------------------------
    def handle_message(self, body) -> bool:
        """What `callback` does between receive and ack: True = basic_ack, False = basic_nack."""
        try:
            message = json.loads(body)
            doc_id = message.get("doc_id")
            text = message.get("original_text_masked")
            chunks = [text[i:i + CHUNK_CHARS] for i in range(0, len(text), CHUNK_CHARS)]
            if self.keep_doc_ids:
                extra = {key: message[key] for key in ("patient_id", "date") if message.get(key) is not None}
                self.add_many(chunks, f"Dossier Patient {doc_id}", "patient_file", doc_id=doc_id, extra=extra)
            else:
                self.add_many(chunks, f"Dossier Patient {doc_id}", "patient_file")
            self.save_state()
            return True
        except Exception as e:
            print(f"Erreur: {e}")
            return False

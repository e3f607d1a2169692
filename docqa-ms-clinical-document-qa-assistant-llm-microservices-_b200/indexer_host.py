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
  * `save_state` is append-only (reference: both files rewritten in full after every message,
    indexer.py:26-30,125 -- O(N) per message).  The rows added since the last save are appended to
    `vector_store.faiss` (only that tail leaves the GPU) and the 45-byte header is rewritten in one
    write; the new metadata records are appended to `metadata_store.pkl`, which is laid out so that it
    stays ONE pickle that `pickle.load` turns into the same `list[dict]` llm-qa/main.py:37-38 expects.
    Both files have a single small write as their commit point, so a reader sees the old state or the
    new one.  The first save, a save after rows were removed, and files this class did not write fall
    back to the full rewrite -- to temporaries renamed into place (reference: two in-place writes, so
    llm-qa can read a torn pair).

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


def _default_append(index, path, rows_on_disk):
    from . import flat

    flat.write_index_append(index, path, rows_on_disk)


# ---- metadata_store.pkl as an appendable pickle ---------------------------------------------------
# pickle.load(f) must keep returning list[dict] (llm-qa/main.py:37-38).  A pickled list is
# `EMPTY_LIST, then any number of (MARK item ... item APPENDS) groups, then STOP`, so records can be added by
# writing one more group.  Records are pickled one by one with protocol 3 (no framing) and the memo switched
# off (no back-references into earlier bytes), which makes their byte strings freely concatenable.
_META_HEAD = b"\x80\x03]"      # PROTO 3, EMPTY_LIST
_MARK, _APPENDS, _STOP = b"(", b"e", b"."


def _record_bytes(record) -> bytes:
    import io

    buf = io.BytesIO()
    pk = pickle.Pickler(buf, protocol=3)
    pk.fast = True  # no memo: the bytes of one record never refer to another
    pk.dump(record)
    raw = buf.getvalue()
    assert raw[:2] == b"\x80\x03" and raw[-1:] == _STOP
    return raw[2:-1]


def _group_bytes(records) -> bytes:
    return (_MARK + b"".join(_record_bytes(r) for r in records) + _APPENDS) if records else b""


def dump_metadata(records, f) -> None:
    """Whole store, in the appendable layout (groups of <= 4096 records)."""
    f.write(_META_HEAD)
    for i in range(0, len(records), 4096):
        f.write(_group_bytes(records[i:i + 4096]))
    f.write(_STOP)


def metadata_is_appendable(path, valid_len: int | None = None) -> bool:
    """True for a file in dump_metadata's layout whose pickle ends at `valid_len` (default: the file's end)."""
    try:
        with open(path, "rb") as f:
            head = f.read(len(_META_HEAD))
            f.seek((valid_len if valid_len is not None else os.path.getsize(path)) - 2)
            tail = f.read(2)
    except (OSError, ValueError):
        return False
    return head == _META_HEAD and tail in (b"]" + _STOP, _APPENDS + _STOP)


def pickle_end(path) -> int:
    """Offset just past the STOP of the pickle at the start of `path` (bytes after it are not part of it)."""
    with open(path, "rb") as f:
        pickle.load(f)
        return f.tell()


def append_metadata(path, records, valid_len: int | None = None) -> int:
    """Add records to a file written by dump_metadata / append_metadata; returns the file's new valid length.

    `valid_len` is where the pickle ends (None: found by reading it).  Whatever follows -- the debris of an
    append that never committed -- is cut off.  The new group is then written AFTER the STOP byte (a reader
    stops at STOP and never sees it), flushed, and only then is that one STOP byte turned into the group's MARK:
    a single-byte write is the commit point."""
    if valid_len is None:
        valid_len = pickle_end(path)
    if not records:
        return valid_len
    group = _group_bytes(records)
    with open(path, "r+b") as f:
        stop_at = valid_len - 1
        f.seek(stop_at)
        if f.read(1) != _STOP:
            raise ValueError(f"{path} does not hold a pickle ending at byte {valid_len}")
        f.truncate(valid_len)
        f.write(group[1:] + _STOP)  # everything but the MARK, which will take the old STOP's place
        f.flush()
        os.fdatasync(f.fileno())
        os.pwrite(f.fileno(), _MARK, stop_at)
        os.fdatasync(f.fileno())
    return valid_len + len(group)


class SemanticIndexer:
    """State and behaviour of the reference's indexer module, as an object."""

    def __init__(self, model, dimension: int = 384, *, index_file: str = INDEX_FILE,
                 metadata_file: str = METADATA_FILE, default_data_dir: str = DEFAULT_DATA_DIR,
                 index_factory=None, write_index=None, read_index=None, append_index=None,
                 keep_doc_ids: bool = False, incremental_save: bool = True, checkpoint_every: int | None = None):
        self.model = model                      # .encode(list[str]) -> [n, d]; optional .encode_tokens
        self.dimension = int(dimension)
        self.index_file = index_file
        self.metadata_file = metadata_file
        self.default_data_dir = default_data_dir
        self._index_factory = index_factory or _default_index_factory
        self._write_index, self._read_index = write_index, read_index  # None -> this package's, resolved lazily
        self._append_index = append_index       # (index, path, rows_on_disk); None -> bfk_write_append
        self.index = None
        self.metadata_store: list[dict] = []
        # append-only persistence: how much of the current state the two files are known to hold
        self.incremental_save = bool(incremental_save)
        self.checkpoint_every = checkpoint_every  # full atomic rewrite every so many saves (None: never)
        self._saved_rows: int | None = None       # rows of self.index in index_file (None: unknown -> full rewrite)
        self._saved_meta: int | None = None       # records in metadata_file, written in the appendable layout
        self._meta_len: int | None = None         # where the pickle in metadata_file ends (None: read it to find out)
        self._saves_since_full = 0
        self.last_save_mode = None                # "full" / "append" / "noop" (what the last save_state did)
        # False: every row carries doc_id "KB_MTC" like the reference (indexer.py:44).  True: patient messages
        # keep their own doc_id plus patient_id / date when the message has them -- what the patient-snippets
        # endpoint (search_service.py; synthese-comparative/core/retrieval_client.py:58-91) filters on.
        self.keep_doc_ids = bool(keep_doc_ids)

    # ---- indexer.py:26-30 ------------------------------------------------------------
    def save_state(self) -> None:
        """Persist index + metadata: append what is new when the files hold a prefix of the state, else rewrite."""
        n = int(self.index.ntotal)
        can_append = (
            self.incremental_save and self._saved_rows is not None and self._saved_meta is not None
            and self._saved_rows <= n and self._saved_meta <= len(self.metadata_store)
            and (self.checkpoint_every is None or self._saves_since_full + 1 < self.checkpoint_every)
            and os.path.exists(self.index_file) and os.path.exists(self.metadata_file)
        )
        if can_append:
            try:
                if self._saved_rows == n and self._saved_meta == len(self.metadata_store):
                    self.last_save_mode = "noop"
                    return
                # metadata first: a reader that sees the new index always finds metadata at least as long
                self._meta_len = append_metadata(self.metadata_file, self.metadata_store[self._saved_meta:], self._meta_len)
                self._saved_meta = len(self.metadata_store)
                if n > self._saved_rows:
                    (self._append_index or _default_append)(self.index, self.index_file, self._saved_rows)
                self._saved_rows = n
                self._saves_since_full += 1
                self.last_save_mode = "append"
                return
            except (RuntimeError, ValueError, OSError) as e:
                print(f"save_state: append refused ({e}); rewriting both files")
        self._save_full()

    def _save_full(self) -> None:
        """Both files written to temporaries and renamed into place (each appears atomically)."""
        d = os.path.dirname(os.path.abspath(self.index_file)) or "."
        fd, tmp_idx = tempfile.mkstemp(prefix=".vector_store.", suffix=".tmp", dir=d)
        os.close(fd)
        fd, tmp_meta = tempfile.mkstemp(prefix=".metadata_store.", suffix=".tmp", dir=d)
        try:
            (self._write_index or _default_io()[0])(self.index, tmp_idx)
            with os.fdopen(fd, "wb") as f:
                dump_metadata(self.metadata_store, f)
            # metadata first: a reader that sees the new index always finds metadata at least as long
            os.replace(tmp_meta, self.metadata_file)
            os.replace(tmp_idx, self.index_file)
        finally:
            for p in (tmp_idx, tmp_meta):
                if os.path.exists(p):
                    os.unlink(p)
        self._saved_rows, self._saved_meta = int(self.index.ntotal), len(self.metadata_store)
        self._meta_len = os.path.getsize(self.metadata_file)
        self._saves_since_full = 0
        self.last_save_mode = "full"

    def invalidate_saved_state(self) -> None:
        """Call after removing or reordering rows (index.remove_ids / reset): the next save rewrites both files."""
        self._saved_rows = self._saved_meta = None

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

    def add_vectors(self, embeddings, records) -> int:
        """Rows that are already embeddings (float32 [n, d]) with their metadata records, appended as
        `index.add` + `metadata_store.append` would (indexer.py:41-48); for callers that encode elsewhere."""
        embeddings = np.asarray(embeddings, dtype="float32").reshape(len(records), -1)
        if len(records) == 0:
            return 0
        if self.index is None:
            self.index = self._index_factory(self.dimension)
        self.index.add(embeddings)
        self.metadata_store.extend(records)
        return len(records)

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
                self._meta_len = f.tell()  # an uncommitted append may have left bytes behind the pickle
            n = int(self.index.ntotal)
            if len(self.metadata_store) > n:
                # a crash between the two commits of an append: those records never became searchable
                self.metadata_store = self.metadata_store[:n]
                self._saved_rows = self._saved_meta = None
            else:
                # the files hold exactly this state; appends are possible if the pickle is in the appendable layout
                # (a file written by the reference's pickle.dump is not: the first save rewrites it)
                ok = metadata_is_appendable(self.metadata_file, self._meta_len)
                self._saved_rows = n if ok else None
                self._saved_meta = len(self.metadata_store) if ok else None
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

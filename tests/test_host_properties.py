"""Property tests (hypothesis) of the round-2 host logic: the appendable metadata pickle (indexer_host.py) and the
block-cyclic id arithmetic of the sharded index (sharded.py).  CPU only, no process group needed."""
import io
import os
import pickle

import numpy as np
from hypothesis import given, settings
from hypothesis import strategies as st

from docqa_b200 import indexer_host
from docqa_b200.sharded import ShardedIndexFlat, _faiss_header, _parse_faiss_header
from oracle import faiss_io

_value = st.one_of(st.text(max_size=40), st.integers(-10**12, 10**12), st.none(), st.booleans(),
                   st.floats(allow_nan=False), st.lists(st.text(max_size=8), max_size=4))
_record = st.dictionaries(st.sampled_from(["doc_id", "text_content", "source", "type", "patient_id", "date", "n"]), _value,
                          min_size=1, max_size=6)


@settings(max_examples=60, deadline=None)
@given(batches=st.lists(st.lists(_record, max_size=5), min_size=1, max_size=6))
def test_appendable_pickle_always_loads_as_the_same_list(tmp_path_factory, batches):
    """Whatever the records and however they are split into appends, pickle.load returns the concatenation
    (llm-qa/main.py:37-38 does nothing but pickle.load), and the commit point is one byte."""
    path = str(tmp_path_factory.mktemp("meta") / "metadata_store.pkl")
    with open(path, "wb") as f:
        indexer_host.dump_metadata(batches[0], f)
    so_far = list(batches[0])
    assert pickle.load(open(path, "rb")) == so_far and indexer_host.metadata_is_appendable(path)
    length = os.path.getsize(path)
    for b in batches[1:]:
        before = open(path, "rb").read()
        new_len = indexer_host.append_metadata(path, b, length)
        after = open(path, "rb").read()
        so_far += b
        assert pickle.load(io.BytesIO(after)) == so_far
        assert new_len == len(after) and indexer_host.pickle_end(path) == new_len
        if b:
            # everything before the old STOP byte is untouched; that byte turned from STOP into MARK
            assert after[: length - 1] == before[: length - 1] and before[length - 1:length] == b"." and after[length - 1:length] == b"("
        length = new_len
    # the reference's own writer produces a file that is NOT appendable (and is then rewritten once)
    with open(path, "wb") as f:
        pickle.dump(so_far, f)
    assert not indexer_host.metadata_is_appendable(path) or so_far == []


class _FakeLocal:
    def __init__(self):
        self.ntotal = 0


def _shard(rank, world, block):
    s = ShardedIndexFlat.__new__(ShardedIndexFlat)  # the id arithmetic needs no process group
    s.rank, s.world, s.block_rows, s.placement, s._global_start, s._ntotal = rank, world, block, "block_cyclic", 0, 0
    s.local = _FakeLocal()
    return s


@settings(max_examples=80, deadline=None)
@given(world=st.integers(1, 8), block=st.integers(1, 50), n=st.integers(0, 700), g0=st.integers(0, 300), span=st.integers(0, 400))
def test_block_cyclic_placement_is_a_bijection(world, block, n, g0, span):
    """Every global row has exactly one owner; local rows are dense, in increasing global order; local -> global ->
    local round-trips; _owned_in tiles any range."""
    shards = [_shard(r, world, block) for r in range(world)]
    owner = (np.arange(n) // block) % world
    seen = np.zeros(n, dtype=int)
    for r, s in enumerate(shards):
        mine = np.nonzero(owner == r)[0]
        s.local.ntotal = len(mine)
        s._ntotal = n
        np.testing.assert_array_equal(s._local_global_ids(), mine)              # dense, ordered
        assert [s._local_of(int(g)) for g in mine] == list(range(len(mine)))    # and back
        runs = s._owned_in(0, n)
        covered = np.concatenate([np.arange(a, b) for a, b in runs]) if runs else np.zeros(0, int)
        np.testing.assert_array_equal(covered, mine)
        seen[covered] += 1
        lo, hi = min(g0, n), min(g0 + span, n)
        sub = s._owned_in(lo, hi)
        got = np.concatenate([np.arange(a, b) for a, b in sub]) if sub else np.zeros(0, int)
        np.testing.assert_array_equal(got, mine[(mine >= lo) & (mine < hi)])
        ids = np.array([-1, *range(len(mine))], dtype=np.int64)
        out = s._to_global(ids)
        assert out[0] == -1 and (out[1:] == mine).all()                         # -1 padding stays -1
    assert (seen == 1).all()


@settings(max_examples=40, deadline=None)
@given(d=st.integers(1, 2048), n=st.integers(0, 3), big=st.integers(0, 2**31 - 1), metric=st.sampled_from([0, 1]))
def test_sharded_header_is_the_oracle_writers_header(d, n, big, metric):
    """The 45 bytes rank 0 writes last are byte for byte what the single-process writer puts in front of n rows."""
    assert _faiss_header(d, n, metric) == faiss_io.write_index_bytes(np.zeros((n, d), np.float32), metric)[:45]
    assert _parse_faiss_header(_faiss_header(d, big, metric)) == (d, big, metric)

"""Parity of the CUDA flat index (through the C ABI) against the CPU oracle. GPU only."""
import ctypes
import os

import numpy as np
import pytest

from oracle import cpu_baseline, faiss_io, flat_ref
from tests.parity import check_knn

pytestmark = pytest.mark.gpu

L2, IP = flat_ref.METRIC_L2, flat_ref.METRIC_INNER_PRODUCT


def unit(rng, n, d):
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x


@pytest.fixture(scope="module")
def bfk():
    import docqa_b200

    assert docqa_b200.get_num_gpus() >= 1, "GPU tests need a CUDA device"
    return docqa_b200


def test_reference_fixture_read_search_write(bfk, golden_dir, fixture_index_bytes, fixture_corpus, tmp_path):
    """BASELINE config 1: the reference's own 649x384 index, nq=1, k=3 and k=5."""
    idx = bfk.read_index(os.path.join(golden_dir, "docqa_vector_store.faiss"))
    assert isinstance(idx, bfk.IndexFlatL2)
    assert (idx.d, idx.ntotal, idx.metric_type, idx.is_trained) == (384, 649, 1, True)
    g = np.load(os.path.join(golden_dir, "docqa_search_golden.npz"))
    q = np.concatenate([fixture_corpus, g["perturbed_queries"]], axis=0)
    for r in range(q.shape[0]):  # one query per call, like llm-qa/main.py:117
        k = 3 if r % 2 else 5
        D, I = idx.search(q[r : r + 1], k)
        check_knn(D, I, q[r : r + 1], fixture_corpus, k, L2, I_ref=g["I"][r : r + 1, :k],
                  D_ref64=g["D64"][r : r + 1, :k], tight=True)
        if r < 649:
            assert D[0, 0] == 0.0  # direct-difference form: duplicates / self are exactly 0
    # all 713 queries in one call (exact path explicitly)
    D, I = idx.search(q, 5, algo="exact")
    check_knn(D, I, q, fixture_corpus, 5, L2, I_ref=g["I"], D_ref64=g["D64"], tight=True)
    # and whatever AUTO picks for a batch
    D, I = idx.search(q, 5)
    check_knn(D, I, q, fixture_corpus, 5, L2, I_ref=g["I"], D_ref64=g["D64"])
    # write_index reproduces the reference artefact byte for byte
    out = tmp_path / "vector_store.faiss"
    bfk.write_index(idx, str(out))
    assert out.read_bytes() == fixture_index_bytes
    assert bfk.serialize_index(idx).tobytes() == fixture_index_bytes
    idx2 = bfk.deserialize_index(fixture_index_bytes)
    assert idx2.ntotal == 649 and (idx2.reconstruct_n(0, 649) == fixture_corpus).all()


def test_incremental_add_like_the_indexer(bfk, fixture_corpus):
    """indexer.py:41 adds one (1, 384) row per call; ids are sequential."""
    idx = bfk.IndexFlatL2(384)
    assert idx.ntotal == 0
    for i in range(40):
        idx.add(np.array(fixture_corpus[i : i + 1]).astype("float32"))
        assert idx.ntotal == i + 1
    idx.add(fixture_corpus[40:].astype(np.float64))  # dtype is cast, as faiss's wrapper does
    assert idx.ntotal == 649
    assert (idx.reconstruct_n(0, 649) == fixture_corpus).all()
    assert (idx.reconstruct(648) == fixture_corpus[648]).all()
    D, I = idx.search(fixture_corpus[100:101], 3)
    Dr, Ir = flat_ref.knn(fixture_corpus[100:101], fixture_corpus, 3, L2)
    check_knn(D, I, fixture_corpus[100:101], fixture_corpus, 3, L2, tight=True)


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d", [384, 768, 100, 3, 1])
@pytest.mark.parametrize("nq", [1, 2, 3, 5, 8, 13])
def test_exact_path_random(bfk, metric, d, nq):
    rng = np.random.default_rng(d * 100 + nq + metric)
    n = 5003
    xb = unit(rng, n, d) if d > 1 else rng.standard_normal((n, 1)).astype(np.float32)
    xq = unit(rng, nq, d) if d > 1 else rng.standard_normal((nq, 1)).astype(np.float32)
    idx = bfk.IndexFlat(d, metric)
    idx.add(xb)
    for k in (1, 10, 100):
        D, I = idx.search(xq, k, algo="exact")
        check_knn(D, I, xq, xb, k, metric, tight=(metric == L2 or d > 1))
        Dc, Ic = cpu_baseline.knn_seq(xq, xb, k, metric)
        assert (I == Ic).mean() > 0.995


@pytest.mark.parametrize("metric", [L2, IP])
def test_edge_cases(bfk, metric):
    pad = np.float32(3.4028235e38) if metric == L2 else np.float32(-3.4028235e38)
    rng = np.random.default_rng(5)
    idx = bfk.IndexFlat(16, metric)
    # empty index
    D, I = idx.search(unit(rng, 3, 16), 4)
    assert D.shape == (3, 4) and (I == -1).all() and (D == pad).all()
    # zero queries
    D, I = idx.search(np.zeros((0, 16), np.float32), 4)
    assert D.shape == (0, 4) and I.shape == (0, 4)
    # k > ntotal
    xb = unit(rng, 5, 16)
    idx.add(xb)
    xq = unit(rng, 2, 16)
    D, I = idx.search(xq, 9)
    check_knn(D, I, xq, xb, 9, metric, tight=True)
    assert (I[:, 5:] == -1).all() and (D[:, 5:] == pad).all()
    # caller-provided outputs
    D2 = np.empty((2, 9), np.float32)
    I2 = np.empty((2, 9), np.int64)
    r = idx.search(xq, 9, D=D2, I=I2)
    assert r[0] is D2 and (I2 == I).all()
    # argument checks are assertions, like faiss's python wrapper
    with pytest.raises(AssertionError):
        idx.search(np.zeros((1, 15), np.float32), 3)
    with pytest.raises(AssertionError):
        idx.search(xq, 0)
    with pytest.raises(AssertionError):
        idx.add(np.zeros((1, 17), np.float32))
    with pytest.raises(RuntimeError):
        idx.search(xq, 5000)  # beyond MAX_K
    with pytest.raises(RuntimeError):
        idx.reconstruct(99)
    with pytest.raises(RuntimeError):
        idx.add_with_ids(xb, np.arange(5))
    # maximum k
    xb2 = unit(rng, 3000, 16)
    idx.reset()
    assert idx.ntotal == 0
    idx.add(xb2)
    D, I = idx.search(xq, 2048)
    check_knn(D, I, xq, xb2, 2048, metric)


def test_duplicates_and_adversarial_order(bfk):
    """1% exact duplicates + rows sorted by decreasing distance (every row improves the top-k)."""
    rng = np.random.default_rng(11)
    d, n = 64, 20000
    xb = unit(rng, n, d)
    dup = rng.integers(0, n, size=200)
    xb[rng.integers(0, n, size=200)] = xb[dup]
    q = unit(rng, 4, d)
    order = np.argsort(-((xb - q[0]) ** 2).sum(1))  # farthest first for query 0
    xb = np.ascontiguousarray(xb[order])
    idx = bfk.IndexFlatL2(d)
    idx.add(xb)
    for k in (1, 10, 64):
        D, I = idx.search(q, k, algo="exact")
        check_knn(D, I, q, xb, k, L2, tight=True)
    # all-identical corpus: any k ids are right, distances all equal, smallest ids win
    same = np.repeat(unit(rng, 1, d), 1000, axis=0)
    idx2 = bfk.IndexFlatL2(d)
    idx2.add(same)
    D, I = idx2.search(q[:1], 10)
    assert I[0].tolist() == list(range(10)) and np.ptp(D) == 0


def test_remove_ids_merge_from_reset(bfk):
    rng = np.random.default_rng(6)
    xb = unit(rng, 300, 32)
    idx = bfk.IndexFlatL2(32)
    ref = flat_ref.IndexFlatRef(32)
    idx.add(xb)
    ref.add(xb)
    ids = np.array([0, 5, 5, 17, 299, 1000, -3], dtype=np.int64)
    assert idx.remove_ids(ids) == ref.remove_ids(ids) == 4
    assert idx.ntotal == ref.ntotal == 296
    assert (idx.reconstruct_n(0, 296) == ref.reconstruct_n(0, 296)).all()
    q = unit(rng, 3, 32)
    D, I = idx.search(q, 7)
    check_knn(D, I, q, ref.reconstruct_n(0, 296), 7, L2, tight=True)
    other = bfk.IndexFlatL2(32)
    other.add(xb[:10])
    idx.merge_from(other)
    assert idx.ntotal == 306 and other.ntotal == 0
    assert (idx.reconstruct_n(296, 10) == xb[:10]).all()


def test_ip_file_round_trip_and_bf16_storage(bfk, tmp_path):
    rng = np.random.default_rng(8)
    xb = unit(rng, 1000, 48)
    idx = bfk.IndexFlatIP(48)
    idx.add(xb)
    p = tmp_path / "ip.faiss"
    bfk.write_index(idx, str(p))
    assert p.read_bytes() == faiss_io.write_index_bytes(xb, IP)
    back = bfk.read_index(str(p))
    assert isinstance(back, bfk.IndexFlatIP) and back.ntotal == 1000
    # bf16-stored corpus: the oracle sees the bf16 -> fp32 upcast (BASELINE config 3)
    import torch

    xb16 = torch.from_numpy(xb).to(torch.bfloat16).to(torch.float32).numpy()
    idx16 = bfk.IndexFlatIP(48, storage="bf16")
    idx16.add(xb)
    assert (idx16.reconstruct_n(0, 1000) == xb16).all()
    q = unit(rng, 5, 48)
    D, I = idx16.search(q, 10)
    check_knn(D, I, q, xb16, 10, IP, tight=True)
    idx16l2 = bfk.IndexFlatL2(48, storage="bf16")
    idx16l2.add(xb)
    D, I = idx16l2.search(q, 10)
    check_knn(D, I, q, xb16, 10, L2, tight=True)


def test_torch_cuda_in_cuda_out(bfk):
    import torch

    rng = np.random.default_rng(9)
    xb, q = unit(rng, 4000, 384), unit(rng, 6, 384)
    idx = bfk.IndexFlatL2(384)
    idx.add(torch.from_numpy(xb).cuda())
    D, I = idx.search(torch.from_numpy(q).cuda(), 10)
    assert D.is_cuda and I.is_cuda and I.dtype == torch.int64
    check_knn(D.cpu().numpy(), I.cpu().numpy(), q, xb, 10, L2, tight=True)
    # id_offset (used by the row-sharded index) shifts ids but not the -1 padding
    D2, I2 = idx.search(q[:1], 10, id_offset=1000)
    assert (I2 == I[:1].cpu().numpy() + 1000).all()


def test_faiss_shim_runs_the_indexer_call_sequence(bfk, fixture_corpus, tmp_path):
    """`import faiss` -> this library; then indexer.py:39-41,27,99 and llm-qa/main.py:35,101."""
    bfk.install_as_faiss(force=True)
    import faiss

    index = faiss.IndexFlatL2(384)
    for i in range(5):
        index.add(np.array(fixture_corpus[i : i + 1]).astype("float32"))
    path = str(tmp_path / "vector_store.faiss")
    faiss.write_index(index, path)
    raw_index = faiss.read_index(path)
    assert raw_index.ntotal == 5
    vector = np.array([fixture_corpus[3].tolist()], dtype=np.float32)  # LangChain's call shape
    scores, indices = raw_index.search(vector, 3)
    assert indices[0][0] == 3 and scores[0][0] == 0.0


def test_full_size_properties_1m_x_384(bfk):
    """BASELINE config 2 at full size: oracle on a query subsample + size-independent properties."""
    import torch

    n, d, k = 1_000_000, 384, 10
    g = torch.Generator(device="cuda").manual_seed(1234)
    xb = torch.randn((n, d), generator=g, device="cuda", dtype=torch.float32)
    xb /= xb.norm(dim=1, keepdim=True)
    g2 = torch.Generator(device="cuda").manual_seed(4321)
    q = torch.randn((64, d), generator=g2, device="cuda", dtype=torch.float32)
    q /= q.norm(dim=1, keepdim=True)
    idx = bfk.IndexFlatL2(d)
    idx.add(xb)
    assert idx.ntotal == n
    D, I = idx.search(q, k)
    D, I = D.cpu().numpy(), I.cpu().numpy()
    xb_h, q_h = xb.cpu().numpy(), q.cpu().numpy()
    # oracle (threaded C restatement) on 8 of the queries
    Dc, Ic = cpu_baseline.knn(q_h[:8], xb_h, k, L2)
    assert (I[:8] == Ic).all()
    np.testing.assert_allclose(D[:8], Dc, rtol=1e-3, atol=1e-6)
    # properties: sorted, in range, unique, distances recomputable from the ids
    assert (np.diff(D, axis=1) >= 0).all() and I.min() >= 0 and I.max() < n
    assert all(len(set(r)) == k for r in I.tolist())
    rec = ((xb_h[I.reshape(-1)].reshape(64, k, d) - q_h[:, None, :]) ** 2).sum(-1)
    np.testing.assert_allclose(D, rec, rtol=1e-3, atol=1e-6)
    # querying with corpus rows returns the row itself at distance 0
    rows = torch.tensor([0, 123456, n - 1], device="cuda")
    D0, I0 = idx.search(xb[rows], 1)
    assert I0.flatten().tolist() == rows.tolist() and (D0 == 0).all()
    # nq=1 and nq=64 agree with each other
    D1, I1 = idx.search(q[:1], k)
    assert (I1.cpu().numpy() == I[:1]).all()


def test_indexer_host_on_the_gpu_index(bfk, tmp_path):
    """SURVEY 8f row 1 end to end: message -> chunks -> one batched add on the GPU -> atomic files -> reload."""
    import hashlib
    import json
    import pickle

    class Model:
        def encode(self, texts):
            out = np.empty((len(texts), 384), np.float32)
            for i, t in enumerate(texts):
                rng = np.random.default_rng(int(hashlib.sha1(t.encode()).hexdigest()[:8], 16))
                v = rng.standard_normal(384)
                out[i] = v / np.linalg.norm(v)
            return out

    ix = bfk.SemanticIndexer(Model(), 384, index_file=str(tmp_path / "vector_store.faiss"),
                             metadata_file=str(tmp_path / "metadata_store.pkl"))
    body = json.dumps({"doc_id": 3, "original_text_masked": "x" * 700 + "y" * 700}).encode()
    assert ix.handle_message(body) is True
    assert isinstance(ix.index, bfk.IndexFlatL2) and ix.index.ntotal == 3
    raw_index = bfk.read_index(str(tmp_path / "vector_store.faiss"))          # llm-qa/main.py:35
    meta = pickle.load(open(tmp_path / "metadata_store.pkl", "rb"))             # llm-qa/main.py:37-38
    assert raw_index.ntotal == len(meta) == 3
    vec = Model().encode([meta[1]["text_content"]])
    D, I = raw_index.search(vec, 3)                                             # llm-qa/main.py:101,117
    assert I[0, 0] == 1 and D[0, 0] == 0.0


@pytest.mark.parametrize("storage", ["f32", "bf16"])
def test_write_index_append_is_byte_identical_to_a_full_write(bfk, tmp_path, storage):
    """bfk_write_append (include/bfk.h): the file grows by the new rows only and equals faiss.write_index's output."""
    rng = np.random.default_rng(31)
    d = 384
    idx = bfk.IndexFlatL2(d, storage=storage)
    path, full = str(tmp_path / "vector_store.faiss"), str(tmp_path / "full.faiss")
    idx.add(rng.standard_normal((649, d)).astype(np.float32))
    bfk.write_index(idx, path)
    on_disk = 649
    for n_new in (1, 1, 3, 700, 5000, 0, 1):  # single rows like indexer.py:41, a batch beyond the pinned buffer, nothing
        if n_new:
            idx.add(rng.standard_normal((n_new, d)).astype(np.float32))
        bfk.write_index_append(idx, path, on_disk)
        on_disk = idx.ntotal
        bfk.write_index(idx, full)
        assert open(path, "rb").read() == open(full, "rb").read()
    back = bfk.read_index(path)
    assert back.ntotal == idx.ntotal
    np.testing.assert_array_equal(back.reconstruct_n(0, back.ntotal), idx.reconstruct_n(0, idx.ntotal))
    # debris past the committed end (an append that crashed before its header write) is overwritten and cut off
    with open(path, "ab") as f:
        f.write(b"\xff" * 10_000)
    idx.add(rng.standard_normal((2, d)).astype(np.float32))
    bfk.write_index_append(idx, path, on_disk)
    bfk.write_index(idx, full)
    assert open(path, "rb").read() == open(full, "rb").read()
    # a file that does not hold a prefix of this index is refused, not patched
    with pytest.raises(RuntimeError, match="does not describe"):
        bfk.write_index_append(idx, path, on_disk)  # stale row count: the header now says ntotal
    other = bfk.IndexFlatIP(d)
    other.add(rng.standard_normal((idx.ntotal, d)).astype(np.float32))
    with pytest.raises(RuntimeError):
        bfk.write_index_append(other, path, idx.ntotal)
    with pytest.raises(RuntimeError):
        bfk.write_index_append(idx, str(tmp_path / "missing.faiss"), 0)


def test_indexer_host_appends_on_the_gpu_index(bfk, tmp_path):
    """The reference's ingestion pattern (one add per chunk, one save per message, indexer.py:37-41,120-125) on
    the GPU index with append-only saves: the files equal a full rewrite after every message."""
    import json
    import pickle

    rng = np.random.default_rng(5)

    class Model:
        def encode(self, texts):
            v = rng.standard_normal((len(texts), 384)).astype(np.float32)
            return v / np.linalg.norm(v, axis=1, keepdims=True)

    ix = bfk.SemanticIndexer(Model(), 384, index_file=str(tmp_path / "vector_store.faiss"),
                             metadata_file=str(tmp_path / "metadata_store.pkl"))
    modes = []
    for m in range(5):
        assert ix.handle_message(json.dumps({"doc_id": m, "original_text_masked": "t" * 1200}).encode()) is True
        modes.append(ix.last_save_mode)
        bfk.write_index(ix.index, str(tmp_path / "full.faiss"))
        assert open(tmp_path / "vector_store.faiss", "rb").read() == open(tmp_path / "full.faiss", "rb").read()
        assert pickle.load(open(tmp_path / "metadata_store.pkl", "rb")) == ix.metadata_store
    assert modes == ["full", "append", "append", "append", "append"] and ix.index.ntotal == 15


def test_host_buffers_pinned_staged_and_large(bfk):
    """The three ways host queries / results travel (csrc/bfk_index.cu search_locked): page-locked buffers the kernels
    address directly, small pageable ones through the index's pinned staging buffer, large pageable ones by copies.
    All must give the answer of the device-resident call, bit for bit."""
    import torch

    rng = np.random.default_rng(41)
    d, n, k = 384, 50_000, 10
    xb = rng.standard_normal((n, d)).astype(np.float32)
    xb /= np.linalg.norm(xb, axis=1, keepdims=True)
    idx = bfk.IndexFlatL2(d)
    idx.add(xb)
    for nq in (1, 64, 700):  # 700 x 384 x 4 B = 1.05 MB: beyond the 512 KB staging half
        q = rng.standard_normal((nq, d)).astype(np.float32)
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        Dd, Id = idx.search(torch.from_numpy(q).cuda(), k, algo="tensor")
        Dd, Id = Dd.cpu().numpy(), Id.cpu().numpy()
        check_knn(Dd, Id, q, xb, k, flat_ref.METRIC_L2, tight=True)
        # pageable in, pageable out
        Dp, Ip = idx.search(q, k, algo="tensor")
        assert (Ip == Id).all() and (Dp == Dd).all()
        # page-locked in and out, twice (the second call finds the buffers holding the first answer)
        qp = torch.empty((nq, d), dtype=torch.float32, pin_memory=True)
        qp.copy_(torch.from_numpy(q))
        Dpin = torch.full((nq, k), -1.0, dtype=torch.float32, pin_memory=True)
        Ipin = torch.full((nq, k), -5, dtype=torch.int64, pin_memory=True)
        for _ in range(2):
            idx.search(qp.numpy(), k, D=Dpin.numpy(), I=Ipin.numpy(), algo="tensor")
            assert (Ipin.numpy() == Id).all() and (Dpin.numpy() == Dd).all()
            Dpin.fill_(-1.0)
            Ipin.fill_(-5)
        # mixed: pinned queries, pageable results
        Dm, Im = idx.search(qp.numpy(), k, algo="tensor")
        assert (Im == Id).all() and (Dm == Dd).all()
    # an index whose every answer needs the fall-back (identical rows) through the pinned route: the rewritten
    # rows must have reached the host buffers when the call returns
    same = np.repeat(xb[:1], 3000, axis=0)
    idx2 = bfk.IndexFlatL2(d)
    idx2.add(same)
    qp = torch.empty((12, d), dtype=torch.float32, pin_memory=True)
    qp.copy_(torch.from_numpy(xb[1:13]))
    Dpin = torch.empty((12, k), dtype=torch.float32, pin_memory=True)
    Ipin = torch.empty((12, k), dtype=torch.int64, pin_memory=True)
    idx2.search(qp.numpy(), k, D=Dpin.numpy(), I=Ipin.numpy(), algo="tensor")
    assert (Ipin.numpy() == np.arange(k)[None, :]).all()


def test_profile_stamps_bracket_the_scan(bfk):
    """bfk_profile_begin / end / stamps (include/bfk.h): per tensor pass, preparation start <= scan start <= scan end
    <= re-rank end on the device clock, and the summed scan time is below the summed pass time."""
    import torch

    rng = np.random.default_rng(43)
    xb = torch.from_numpy(rng.standard_normal((200_000, 384)).astype(np.float32)).cuda()
    idx = bfk.IndexFlatL2(384)
    idx.add(xb)
    q = torch.from_numpy(rng.standard_normal((64, 384)).astype(np.float32)).cuda()
    idx.search(q, 10, algo="tensor")
    idx.profile_begin()
    for _ in range(7):
        idx.search(q, 10, algo="tensor")
    prof = idx.profile_end(detail=True)
    ts = idx.profile_stamps().astype(np.int64)
    assert prof["n"] == 7 and ts.shape == (7, 4)
    assert (ts[:, 2] <= ts[:, 0]).all() and (ts[:, 0] < ts[:, 1]).all() and (ts[:, 1] <= ts[:, 3]).all()
    assert (np.diff(ts[:, 2]) > 0).all()  # passes in order
    assert 0 < prof["scan_ms"] <= prof["pass_ms"] and prof["idle_ms"] >= 0
    np.testing.assert_allclose(prof["scan_ms"], (ts[:, 1] - ts[:, 0]).sum() * 1e-6, rtol=1e-9)
    # unprofiled searches leave no stamps behind; end without begin is an error, not a crash
    idx.search(q, 10, algo="tensor")
    with pytest.raises(RuntimeError):
        idx.profile_end()


def test_storage_grows_in_place(bfk):
    """index.add in many steps (the reference appends one row per call, indexer.py:41): with the growable
    virtual-memory storage nothing is copied and no second copy of the corpus is ever resident, so the device
    memory in use follows the rows stored (2 308 B per row at d = 384: fp32 row + bf16 scan row + norm)."""
    import torch

    rng = np.random.default_rng(47)
    d, step, steps = 384, 100_000, 12
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    free0, _ = torch.cuda.mem_get_info()
    idx = bfk.IndexFlatL2(d)
    peak_used = 0
    chunks = []
    for s in range(steps):
        x = rng.standard_normal((step, d)).astype(np.float32)
        chunks.append(x[:50].copy())
        idx.add(x)
        free_now, _ = torch.cuda.mem_get_info()
        peak_used = max(peak_used, free0 - free_now)
    n = step * steps
    assert idx.ntotal == n
    need = n * (d * 4 + d * 2 + 4)
    # growth adds a quarter at a time plus mapping granularity: well under the 2x (+ old copy) of realloc-and-copy
    assert peak_used <= 1.45 * need + (64 << 20), (peak_used, need)
    for s in (0, 5, steps - 1):  # rows written before a growth step are still where they were
        np.testing.assert_array_equal(idx.reconstruct_n(s * step, 50), chunks[s])
    q = chunks[3][:4]
    D, I = idx.search(q, 1)
    assert I[:, 0].tolist() == [3 * step + i for i in range(4)] and (D[:, 0] == 0).all()
    # reserve maps up to the requested rows; removing rows moves the survivors into fresh arrays
    idx.reserve(n + 500_000)
    assert idx.remove_ids(np.arange(0, 1000)) == 1000
    np.testing.assert_array_equal(idx.reconstruct_n(0, 50), x_first_after_removal(chunks, step))
    del idx
    torch.cuda.synchronize()
    free_end, _ = torch.cuda.mem_get_info()
    assert free0 - free_end < (96 << 20)  # everything mapped was given back


def x_first_after_removal(chunks, step):
    # rows 1000.. of the first add step are not in `chunks` (only its first 50 rows are kept): rebuild them
    rng = np.random.default_rng(47)
    x = rng.standard_normal((step, 384)).astype(np.float32)
    return x[1000:1050]


def test_cudamalloc_storage_fallback_in_a_fresh_process():
    """BFK_VMEM=0 (or a driver without virtual memory management): the round-1 storage, cudaMalloc + copy on growth."""
    import subprocess
    import sys

    code = (
        "import numpy as np, docqa_b200 as bfk\n"
        "rng = np.random.default_rng(1)\n"
        "idx = bfk.IndexFlatL2(64)\n"
        "xs = [rng.standard_normal((3000, 64)).astype(np.float32) for _ in range(5)]\n"
        "for x in xs: idx.add(x)\n"
        "xb = np.concatenate(xs)\n"
        "assert np.array_equal(idx.reconstruct_n(0, 15000), xb)\n"
        "D, I = idx.search(xb[7000:7003], 1)\n"
        "assert I[:, 0].tolist() == [7000, 7001, 7002]\n"
        "assert idx.remove_ids(np.array([0, 1])) == 2 and idx.ntotal == 14998\n"
        "print('fallback ok')\n"
    )
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=300,
                         env={**os.environ, "BFK_VMEM": "0", "PYTHONPATH": root})
    assert out.returncode == 0 and "fallback ok" in out.stdout, out.stderr[-2000:]


def test_streams_appends_and_searches_on_different_torch_streams(bfk, tmp_path):
    """add() with device rows only enqueues the append on the caller's stream (torch streams are non-blocking: the
    legacy stream does not wait for them).  Readers on other streams -- a search, reconstruct / write_index, a
    regrowth -- must still see the rows: the library orders them behind the append with an event."""
    import torch

    rng = np.random.default_rng(53)
    d = 384
    side, other = torch.cuda.Stream(), torch.cuda.Stream()
    idx = bfk.IndexFlatL2(d)
    host = []
    for step in range(6):
        x = rng.standard_normal((30_000, d)).astype(np.float32)
        host.append(x)
        with torch.cuda.stream(side):
            xd = torch.from_numpy(x).to("cuda", non_blocking=False)
            big = torch.randn((4096, 4096), device="cuda")
            for _ in range(3):
                big = big @ big  # keeps `side` busy so that the append really is still pending afterwards
            big = big.clamp(-1, 1)
            idx.add(xd)  # enqueued behind the matmuls
        n = idx.ntotal
        # immediately, from the host (legacy stream) and from another torch stream
        got = idx.reconstruct_n(n - 100, 100)
        np.testing.assert_array_equal(got, x[-100:])
        with torch.cuda.stream(other):
            q = torch.from_numpy(x[-5:]).to("cuda")
            D, I = idx.search(q, 1)
            other.synchronize()
        assert I[:, 0].tolist() == list(range(n - 5, n)) and (D[:, 0] == 0).all()
        D2, I2 = idx.search(x[:3], 1)  # numpy in, legacy stream
        assert I2[:, 0].tolist() == list(range(n - 30_000, n - 30_000 + 3))
    xb = np.concatenate(host)
    path = str(tmp_path / "streams.faiss")
    bfk.write_index(idx, path)
    assert open(path, "rb").read() == faiss_io.write_index_bytes(xb, 1)
    # the workspace is shared by searches on both streams: alternate them without host waits in between
    qs = torch.from_numpy(xb[:64]).to("cuda")
    outs = []
    for i in range(6):
        st = side if i % 2 == 0 else other
        with torch.cuda.stream(st):
            outs.append(idx.search(qs, 3, algo="tensor"))
    torch.cuda.synchronize()
    for D, I in outs:
        assert I[:, 0].tolist() == list(range(64))

"""bfk_search_noalloc: the exact search for callers that cannot allocate device memory.

Some users of this library run where cudaMalloc is off limits once they are serving -- the
process handed all of HBM to its own pool at start-up, or the calling thread must never block
on the driver's allocator.  For them the library keeps the scan's scratch (the per-CTA
partial top-k lists) in a fixed array inside the library image instead of a cudaMalloc'ed
workspace.  This test drives that entry point through the C ABI, twice in a row on the same
array with different shapes, and demands the very same answer the allocating entry point
gives (they run the same kernels, so ids must be identical and distances equal), checks it
against the CPU oracle as well, and checks with cudaMemGetInfo that a warmed call leaves the
amount of free device memory exactly where it was, for every kind of shape the entry point takes.

Collected last on purpose (file name): a failure here cannot hide the other GPU tests.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _unit(rng, n, d):
    x = rng.standard_normal((n, d)).astype(np.float32)
    return x / np.linalg.norm(x, axis=1, keepdims=True)


def _both(bfk, torch, idx, q, k):
    """(D, I) from the existing exact entry point and from the no-alloc one, as numpy."""
    from docqa_b200 import _lib

    L = _lib.load()
    q_dev = torch.from_numpy(q).cuda()
    D_ref, I_ref = idx.search(q_dev, k, algo="exact")
    nq = q.shape[0]
    D = torch.full((nq, k), -7.0, dtype=torch.float32, device="cuda")
    I = torch.full((nq, k), -7, dtype=torch.int64, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.bfk_search_noalloc(idx._h, ctypes.c_void_p(q_dev.data_ptr()), nq, k, ctypes.c_void_p(D.data_ptr()),
                              ctypes.c_void_p(I.data_ptr()), stream)
    assert rc == 0, L.bfk_last_error()
    torch.cuda.synchronize()
    return D_ref.cpu().numpy(), I_ref.cpu().numpy(), D.cpu().numpy(), I.cpu().numpy()


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    import torch

    import docqa_b200 as bfk
    from docqa_b200 import _lib

    assert bfk.get_num_gpus() >= 1
    L = _lib.load()
    rng = np.random.default_rng(2024)
    xb = _unit(rng, 30011, 384)
    idx = bfk.IndexFlatL2(384)
    idx.add(xb)

    # call 1: 8 queries, k = 100 -> fills a large part of the shared array
    q1 = _unit(rng, 8, 384)
    D_ref, I_ref, D, I = _both(bfk, torch, idx, q1, 100)
    assert (I == I_ref).all()
    np.testing.assert_allclose(D, D_ref, rtol=1e-6, atol=1e-7)

    # call 2, straight after, different shape: stale lists of call 1 must not leak in
    q2 = _unit(rng, 3, 384)
    D_ref, I_ref, D, I = _both(bfk, torch, idx, q2, 5)
    assert (I == I_ref).all()
    np.testing.assert_allclose(D, D_ref, rtol=1e-6, atol=1e-7)

    # an inner-product index with an odd dimension (row stride != 16-byte multiple) and k > ntotal
    xb3 = _unit(rng, 50, 37)
    idx3 = bfk.IndexFlatIP(37)
    idx3.add(xb3)
    q3 = _unit(rng, 2, 37)
    D_ref, I_ref, D, I = _both(bfk, torch, idx3, q3, 64)
    assert (I == I_ref).all() and (I[:, 50:] == -1).all()
    np.testing.assert_allclose(D, D_ref, rtol=1e-6, atol=1e-7)

    # sizes beyond the fixed scratch are refused with a message, not silently truncated
    q9 = torch.zeros((9, 384), device="cuda")
    D9 = torch.empty((9, 4), device="cuda")
    I9 = torch.empty((9, 4), dtype=torch.int64, device="cuda")
    rc = L.bfk_search_noalloc(idx._h, ctypes.c_void_p(q9.data_ptr()), 9, 4, ctypes.c_void_p(D9.data_ptr()),
                              ctypes.c_void_p(I9.data_ptr()), None)
    assert rc != 0 and b"at most 8 queries" in L.bfk_last_error()
    rc = L.bfk_search_noalloc(idx._h, ctypes.c_void_p(q9.data_ptr()), 1, 129, ctypes.c_void_p(D9.data_ptr()),
                              ctypes.c_void_p(I9.data_ptr()), None)
    assert rc != 0 and b"k is limited to 128" in L.bfk_last_error()


@pytest.mark.parametrize("storage", ["f32", "bf16"])
@pytest.mark.parametrize("metric", [1, 0])
def test_noalloc_answers_match_the_oracle_and_free_memory_does_not_move(metric, storage):
    """Every (query-count bucket, metric, storage) the entry point accepts: right answer, no device memory taken.

    The first call of a shape may still make the driver load that kernel variant (a one-time cost a caller
    pays at start-up by running each shape once); from the second call on, cudaMemGetInfo must report the
    same number of free bytes before and after -- which is what "allocates nothing" means to a process whose
    memory pool owns the rest of HBM.
    """
    import torch

    import docqa_b200 as bfk
    from docqa_b200 import _lib
    from oracle import flat_ref
    from tests.parity import check_knn

    L = _lib.load()
    rng = np.random.default_rng(77 + metric)
    d, n, k = 96, 20000, 10
    xb = _unit(rng, n, d)
    idx = bfk.IndexFlat(d, metric, storage=storage)
    idx.add(xb)
    xb_stored = idx.reconstruct_n(0, n)  # what the index really holds (bf16-rounded rows for storage="bf16")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for nq in (1, 2, 3, 4, 5, 8):
        q = _unit(rng, nq, d)
        q_dev = torch.from_numpy(q).cuda()
        D = torch.empty((nq, k), dtype=torch.float32, device="cuda")
        I = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        args = (idx._h, ctypes.c_void_p(q_dev.data_ptr()), nq, k, ctypes.c_void_p(D.data_ptr()),
                ctypes.c_void_p(I.data_ptr()), stream)
        assert L.bfk_search_noalloc(*args) == 0, L.bfk_last_error()  # warms this shape
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        D.fill_(-7.0)
        I.fill_(-7)
        assert L.bfk_search_noalloc(*args) == 0, L.bfk_last_error()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after == free_before, (nq, metric, storage, free_before - free_after)
        check_knn(D.cpu().numpy(), I.cpu().numpy(), q, xb_stored, k, metric)

"""Row-sharded host logic on CPU: world_size 2, gloo backend (not gpu).

The local searcher and the merge are injected from the oracle here (the product defaults
are the CUDA library); what is under test is the sharding arithmetic, global ids, the
single all-gather and its rank ordering, the block-cyclic placement of replicated adds, and the
drop-in surface on the sharded object (write_index / read_index / reconstruct / remove_ids /
selector search / range_search).
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import flat_ref


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def oracle_merge(Dg, Ig, k, metric):
    W, nq, _ = Dg.shape
    D = Dg.permute(1, 0, 2).reshape(nq, W * k).numpy()
    I = Ig.permute(1, 0, 2).reshape(nq, W * k).numpy()
    key = np.where(I >= 0, D if metric == flat_ref.METRIC_L2 else -D, np.inf)
    outD = np.empty((nq, k), np.float32)
    outI = np.empty((nq, k), np.int64)
    for r in range(nq):
        o = np.lexsort((np.arange(W * k), key[r]))[:k]
        outD[r], outI[r] = D[r, o], I[r, o]
    return torch.from_numpy(outD), torch.from_numpy(outI)


def _worker(rank, world, port, metric, ret, tmp):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from docqa_b200 import IDSelectorRange, SearchParameters, ShardedIndexFlat, shard_range
        from oracle import faiss_io

        def mk(**kw):
            return ShardedIndexFlat(32, metric, local_index=flat_ref.IndexFlatRef(32, metric), merge_fn=oracle_merge, **kw)

        rng = np.random.default_rng(7)
        xb = rng.standard_normal((501, 32)).astype(np.float32)
        xq = rng.standard_normal((9, 32)).astype(np.float32)
        # replicated adds, the reference's call shape included (one row per call, indexer.py:41): blocks of 64 rows
        # go round the ranks, so both ranks fill up and ids stay arithmetic
        idx = mk(block_rows=64)
        idx.add(xb[:300])
        for i in range(300, 330):
            idx.add(xb[i:i + 1])
        idx.add(xb[330:])
        assert idx.ntotal == 501
        owned = [g for g in range(501) if (g // 64) % world == rank]
        assert idx.local.ntotal == len(owned)
        np.testing.assert_array_equal(idx._local_global_ids(), np.array(owned))
        np.testing.assert_array_equal(idx.local.reconstruct_n(0, len(owned)), xb[owned])
        D, I = idx.search(xq, 10)
        Dr, Ir = flat_ref.knn(xq, xb, 10, metric)
        assert (I == Ir).all(), (I, Ir)
        np.testing.assert_allclose(D, Dr, rtol=1e-5, atol=1e-6)
        # reconstruct / reconstruct_n: every rank sees every row
        np.testing.assert_array_equal(idx.reconstruct_n(0, 501), xb)
        np.testing.assert_array_equal(idx.reconstruct_n(60, 200), xb[60:260])
        np.testing.assert_array_equal(idx.reconstruct(333), xb[333])
        # write_index: one faiss-format file, byte-identical to the single-process writer (indexer.py:27)
        path = os.path.join(tmp, f"sharded_{metric}.faiss")
        idx.write_index(path)
        assert open(path, "rb").read() == faiss_io.write_index_bytes(xb, metric)
        # read_index onto the ranks (indexer.py:99, llm-qa/main.py:35), both placements; same answers
        for placement in ("block_cyclic", "contiguous"):
            back = ShardedIndexFlat.read_index(path, local_index_factory=lambda d, m: flat_ref.IndexFlatRef(d, m),
                                               merge_fn=oracle_merge, placement=placement, block_rows=64)
            assert back.ntotal == 501 and back.d == 32 and back.metric_type == metric
            Db, Ib = back.search(xq, 10)
            assert (Ib == Ir).all()
            np.testing.assert_array_equal(back.reconstruct_n(0, 501), xb)
        # search restricted by a selector over GLOBAL ids (faiss SearchParameters)
        Ds, Is = idx.search(xq, 5, params=SearchParameters(sel=IDSelectorRange(100, 400)))
        Drs, Irs = flat_ref.knn(xq, xb[100:400], 5, metric)
        assert (Is == Irs + 100).all()
        # range_search with global ids in id order
        thr = float(np.median(Dr[:, 5]))
        lims, Dg, Ig = idx.range_search(xq, thr)
        ref = flat_ref.IndexFlatRef(32, metric)
        ref.add(xb)
        lr, Drr, Irr = ref.range_search(xq, thr)
        np.testing.assert_array_equal(lims, lr)
        for q in range(9):
            a, b = lims[q], lims[q + 1]
            o = np.argsort(Irr[a:b], kind="stable")
            np.testing.assert_array_equal(Ig[a:b], Irr[a:b][o])
        # remove_ids: survivors keep their order and are renumbered; rows are re-dealt to their new owners
        gone = np.array([0, 5, 63, 64, 200, 499, 9999])
        assert idx.remove_ids(gone) == 6
        keep = np.setdiff1d(np.arange(501), gone)
        assert idx.ntotal == 495
        np.testing.assert_array_equal(idx.reconstruct_n(0, 495), xb[keep])
        owned = [g for g in range(495) if (g // 64) % world == rank]
        assert idx.local.ntotal == len(owned)
        D4, I4 = idx.search(xq, 10)
        Dr4, Ir4 = flat_ref.knn(xq, xb[keep], 10, metric)
        assert (I4 == Ir4).all()
        idx.add(xb[:7])  # and the index keeps growing in place afterwards
        assert idx.ntotal == 502
        np.testing.assert_array_equal(idx.reconstruct_n(495, 7), xb[:7])
        # rank-local adds with explicit global offsets (the benchmark's layout)
        idx2 = mk()
        lo, hi = shard_range(501, rank, world)
        idx2.add_local(xb[lo:hi], lo, 501)
        assert idx2.ntotal == 501
        D2, I2 = idx2.search(torch.from_numpy(xq), 10)
        assert (I2.numpy() == Ir).all()
        np.testing.assert_array_equal(idx2.reconstruct_n(0, 501), xb)
        path2 = os.path.join(tmp, f"contig_{metric}.faiss")
        idx2.write_index(path2)
        assert open(path2, "rb").read() == faiss_io.write_index_bytes(xb, metric)
        with pytest.raises(RuntimeError):
            idx2.add(xb[:3])  # placements cannot be mixed
        # k larger than a shard: padding must survive the merge
        D3, I3 = idx2.search(xq[:2], 400)
        Dr3, Ir3 = flat_ref.knn(xq[:2], xb, 400, metric)
        assert (I3 == Ir3).all()
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("metric", [flat_ref.METRIC_L2, flat_ref.METRIC_INNER_PRODUCT])
def test_sharded_search_world2_gloo(metric, tmp_path):
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, metric, ret, str(tmp_path)), nprocs=2, join=True)
    assert dict(ret) == {0: "ok", 1: "ok"}


def test_shard_range_covers_everything():
    from docqa_b200 import shard_range

    for n in (0, 1, 7, 100, 100_000_000):
        for w in (1, 2, 3, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1

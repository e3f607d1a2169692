"""Row-sharded host logic on CPU: world_size 2, gloo backend (not gpu).

The local searcher and the merge are injected from the oracle here (the product defaults
are the CUDA library); what is under test is the sharding arithmetic, global ids, the
single all-gather and its rank ordering.
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


def _worker(rank, world, port, metric, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from docqa_b200 import ShardedIndexFlat, shard_range

        rng = np.random.default_rng(7)
        xb = rng.standard_normal((501, 32)).astype(np.float32)
        xq = rng.standard_normal((9, 32)).astype(np.float32)
        idx = ShardedIndexFlat(32, metric, local_index=flat_ref.IndexFlatRef(32, metric), merge_fn=oracle_merge)
        # replicated adds in two batches -> two id segments per rank
        idx.add(xb[:300])
        idx.add(xb[300:])
        assert idx.ntotal == 501
        lo, hi = shard_range(300, rank, world)
        assert idx.local.ntotal == (hi - lo) + (shard_range(201, rank, world)[1] - shard_range(201, rank, world)[0])
        D, I = idx.search(xq, 10)
        Dr, Ir = flat_ref.knn(xq, xb, 10, metric)
        assert (I == Ir).all(), (I, Ir)
        np.testing.assert_allclose(D, Dr, rtol=1e-5, atol=1e-6)
        # rank-local adds with explicit global offsets (the benchmark's layout)
        idx2 = ShardedIndexFlat(32, metric, local_index=flat_ref.IndexFlatRef(32, metric), merge_fn=oracle_merge)
        lo, hi = shard_range(501, rank, world)
        idx2.add_local(xb[lo:hi], lo, 501)
        assert idx2.ntotal == 501
        D2, I2 = idx2.search(torch.from_numpy(xq), 10)
        assert (I2.numpy() == Ir).all()
        # k larger than a shard: padding must survive the merge
        D3, I3 = idx2.search(xq[:2], 400)
        Dr3, Ir3 = flat_ref.knn(xq[:2], xb, 400, metric)
        assert (I3 == Ir3).all()
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("metric", [flat_ref.METRIC_L2, flat_ref.METRIC_INNER_PRODUCT])
def test_sharded_search_world2_gloo(metric):
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, metric, ret), nprocs=2, join=True)
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

"""Row-sharded index on real GPUs: NCCL all-gather of per-GPU top-k + K6 merge. GPU only.

Uses as many GPUs as the box shows (up to 4); with a single GPU it still runs the class
end to end (world size 1) and the merge kernel through the C ABI on synthetic shard lists.
"""
import os
import socket

import numpy as np
import pytest

from oracle import flat_ref
from tests.parity import check_knn

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret, tmp):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import docqa_b200 as bfk

        rng = np.random.default_rng(99)
        xb = rng.standard_normal((60001, 384)).astype(np.float32)
        xb /= np.linalg.norm(xb, axis=1, keepdims=True)
        xq = rng.standard_normal((70, 384)).astype(np.float32)
        xq /= np.linalg.norm(xq, axis=1, keepdims=True)
        for metric in (bfk.METRIC_L2, bfk.METRIC_INNER_PRODUCT):
            # replicated adds, block-cyclic over the ranks (blocks of 4096 rows here), one-row adds included
            idx = bfk.ShardedIndexFlat(384, metric, device=rank, block_rows=4096)
            idx.add(xb[:25000])
            for i in range(25000, 25010):
                idx.add(xb[i:i + 1])  # the reference's call shape, indexer.py:41
            idx.add(xb[25010:])
            assert idx.ntotal == 60001
            assert idx.local.ntotal == sum(min(60001, (b + 1) * 4096) - b * 4096 for b in range(rank, 15, world))
            single = bfk.IndexFlat(384, metric, device=rank)  # the same rows on one GPU: the answers must be identical
            single.add(xb)
            for nq in (1, 5, 70):  # exact path, and the tensor path on each shard; host buffers in and out
                D, I = idx.search(xq[:nq], 10)
                check_knn(D, I, xq[:nq], xb, 10, metric, tight=True)
                Ds, Is = single.search(xq[:nq], 10)
                assert (I == Is).all() and (D == Ds).all()  # bit for bit: the same fp32 re-score on either side
            # one faiss-format file written by all ranks, byte-identical to the single-GPU writer (indexer.py:27) ...
            path = os.path.join(tmp, f"sharded_{metric}.faiss")
            one = os.path.join(tmp, f"single_{metric}_{rank}.faiss")
            idx.write_index(path)
            bfk.write_index(single, one)
            assert open(path, "rb").read() == open(one, "rb").read()
            # ... and read back onto the GPUs (indexer.py:99, llm-qa/main.py:35), both placements
            for placement in ("block_cyclic", "contiguous"):
                back = bfk.ShardedIndexFlat.read_index(path, device=rank, placement=placement, block_rows=4096)
                assert back.ntotal == 60001 and back.metric_type == metric
                Db, Ib = back.search(xq, 10)
                Ds, Is = single.search(xq, 10)
                assert (Ib == Is).all() and (Db == Ds).all()
            np.testing.assert_array_equal(idx.reconstruct_n(4000, 200), xb[4000:4200])
            # selector over global ids, and rows removed: survivors renumbered and re-dealt over the GPUs
            Dsel, Isel = idx.search(xq[:9], 5, params=bfk.SearchParameters(sel=bfk.IDSelectorRange(1000, 30000)))
            check_knn(Dsel, Isel - 1000, xq[:9], xb[1000:30000], 5, metric)
            gone = np.arange(0, 60001, 7)
            assert idx.remove_ids(gone) == gone.size
            keep = np.setdiff1d(np.arange(60001), gone)
            Dk, Ik = idx.search(xq[:20], 10)
            check_knn(Dk, Ik, xq[:20], xb[keep], 10, metric, tight=True)
            # contiguous shards with ids global straight from the kernel, CUDA tensors in and out
            idx2 = bfk.ShardedIndexFlat(384, metric, device=rank, storage="bf16")
            lo, hi = bfk.shard_range(60001, rank, world)
            idx2.add_local(torch.from_numpy(xb[lo:hi]).cuda(), lo, 60001)
            D2, I2 = idx2.search(torch.from_numpy(xq).cuda(), 10)
            xb16 = torch.from_numpy(xb).to(torch.bfloat16).to(torch.float32).numpy()
            check_knn(D2.cpu().numpy(), I2.cpu().numpy(), xq, xb16, 10, metric, tight=True)
            # every rank holds the same answer
            ref = I2.clone()
            dist.broadcast(ref, src=0)
            assert torch.equal(ref, I2)
            # k larger than a shard's share of good rows: padding survives the merge
            idx3 = bfk.ShardedIndexFlat(384, metric, device=rank, block_rows=4096)
            idx3.add(xb[:9000])
            D3, I3 = idx3.search(xq[:3], 1000)
            check_knn(D3, I3, xq[:3], xb[:9000], 1000, metric)
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_sharded_search_nccl(tmp_path):
    import torch
    import torch.multiprocessing as mp

    from tests.conftest import note

    world = max(1, min(4, torch.cuda.device_count()))
    note(f"test_sharded_search_nccl: {torch.cuda.device_count()} GPUs visible, world size {world} "
         f"({'peer-memory exchange K6x + deferred search exercised' if world > 1 else 'single rank: K6x NOT exercised'})")
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret, str(tmp_path)), nprocs=world, join=True)
    assert dict(ret) == {r: "ok" for r in range(world)}


def test_merge_kernel_matches_oracle_merge():
    import torch

    import docqa_b200 as bfk
    from docqa_b200.sharded import _cuda_merge

    rng = np.random.default_rng(5)
    W, nq, k = 8, 33, 10
    for metric in (bfk.METRIC_L2, bfk.METRIC_INNER_PRODUCT):
        D = np.sort(rng.random((W, nq, k)).astype(np.float32), axis=2)
        if metric == bfk.METRIC_INNER_PRODUCT:
            D = D[:, :, ::-1].copy()
        I = rng.permutation(W * nq * k).reshape(W, nq, k).astype(np.int64)
        D[3, :, 7:] = np.float32(3.4028235e38) if metric == bfk.METRIC_L2 else np.float32(-3.4028235e38)
        I[3, :, 7:] = -1  # a shard that ran out of rows
        D[:, 5, 0] = D[0, 5, 0]  # exact ties across shards: lower shard wins
        Dg, Ig = torch.from_numpy(D).cuda(), torch.from_numpy(I).cuda()
        Dm, Im = _cuda_merge(Dg, Ig, k, metric)
        Dm, Im = Dm.cpu().numpy(), Im.cpu().numpy()
        for r in range(nq):
            key = np.where(I[:, r].reshape(-1) >= 0, (D[:, r] if metric == bfk.METRIC_L2 else -D[:, r]).reshape(-1), np.inf)
            o = np.lexsort((np.arange(W * k), key))[:k]
            assert (Im[r] == I[:, r].reshape(-1)[o]).all()
            assert (Dm[r] == D[:, r].reshape(-1)[o]).all()

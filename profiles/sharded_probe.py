"""Where the per-step time of the row-sharded search goes (run under torchrun)."""
import os, sys, time, torch, torch.distributed as dist
sys.path.insert(0, '/root/repo')
import docqa_b200 as bfk
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
n_total, d, k = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000, 384, 10
lo, hi = bfk.shard_range(n_total, rank, world)
idx = bfk.ShardedIndexFlat(d, bfk.METRIC_L2, device=lr, storage="bf16")
idx.local.reserve(hi - lo)
g = torch.Generator(device=dev).manual_seed(1234 + rank)
for r0 in range(lo, hi, 1 << 20):
    m = min(1 << 20, hi - r0)
    x = torch.randn((m, d), generator=g, device=dev); x /= x.norm(dim=1, keepdim=True)
    idx.local.add(x)
idx._record(hi - lo, lo, n_total)
def timeit(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
for nq in (1, 64):
    q = torch.randn((nq, d), device=dev); q /= q.norm(dim=1, keepdim=True)
    t_local = timeit(lambda: idx.local.search(q, k, id_offset=lo))
    t_full = timeit(lambda: idx.search(q, k))
    t_np = timeit(lambda: idx.search(q.cpu().numpy(), k), 10)
    Dl, Il = idx.local.search(q, k, id_offset=lo)
    def pack():
        return torch.stack([Dl.contiguous().view(torch.int32).to(torch.int64), Il.contiguous()], dim=0)
    t_pack = timeit(pack)
    payload = pack(); flat = torch.empty((world * 2, nq, k), dtype=torch.int64, device=dev)
    t_ag = timeit(lambda: dist.all_gather_into_tensor(flat, payload))
    gathered = flat.view(world, 2, nq, k)
    def unpack():
        return gathered[:, 0].to(torch.int32).view(torch.float32).contiguous(), gathered[:, 1].contiguous()
    t_un = timeit(unpack)
    Dg, Ig = unpack()
    from docqa_b200.sharded import _cuda_merge
    t_mg = timeit(lambda: _cuda_merge(Dg, Ig, k, 1))
    # the peer-memory exchange alone (static lists), and the deferred local search + finish alone
    import ctypes
    from docqa_b200 import _lib
    L = _lib.load()
    xh = idx._exchange(nq * k, dev)
    b = idx._gpu_buffers(nq, k, dev)
    Do = torch.empty((nq, k), device=dev); Io = torch.empty((nq, k), dtype=torch.int64, device=dev)
    stp = ctypes.c_void_p(torch.cuda.current_stream(dev.index).cuda_stream)
    def xonly():
        _lib.check(L.bfk_xchg_merge(xh, ctypes.c_void_p(b["D"].data_ptr()), ctypes.c_void_p(b["I"].data_ptr()), nq, k, 1,
                                    None, ctypes.c_void_p(Do.data_ptr()), ctypes.c_void_p(Io.data_ptr()), stp))
    t_x = timeit(xonly)
    def xsync():
        xonly(); torch.cuda.current_stream().synchronize()
    t_xs = timeit(xsync)
    def defer_finish():
        idx.local.search(q, k, D=b["D"], I=b["I"], id_offset=lo, defer=True); idx.local.finish()
    t_df = timeit(defer_finish)
    if rank == 0:
        print(f"nq={nq} xchg pipelined {t_x:.3f} xchg+sync {t_xs:.3f} defer+finish {t_df:.3f}", flush=True)
        print(f"nq={nq} world={world}: local {t_local:.3f} ms | full {t_full:.3f} numpy-api {t_np:.3f} | pack {t_pack:.3f} allgather {t_ag:.3f} unpack {t_un:.3f} merge {t_mg:.3f}", flush=True)
dist.destroy_process_group()

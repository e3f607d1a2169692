// K6x: exchange of per-GPU top-k lists over NVLink peer memory, fused with the merge.
//
// Replaces the NCCL all-gather of the row-sharded search (BASELINE config 5; no reference counterpart --
// the reference is a single CPU process) for the latency-bound regime: at nq <= 64 the per-GPU scan of a
// 100M-row corpus split 8 ways takes ~1.4 ms, and two host-launched collectives per search (the gather and
// the "did any rank re-run a query" agreement) cost ~0.8 ms of rank skew and launch latency on top.
// Here every rank owns one exchange region in its HBM, mapped into all peers through CUDA IPC:
//
//   region = 2 parities x [ W slots x slot_bytes | W flags ]
//
//   push kernel   : rank r stores its packed (D, I) lists and its count of uncertified queries into
//                   slot r of EVERY rank's region (plain st.global on peer pointers, through NVSwitch),
//                   then __threadfence_system() and a release-store of the search's sequence number to
//                   flag r of every region;
//   wait kernel   : one warp acquires the W local flags (lane = rank) until all carry the sequence number,
//                   with a clock watchdog (a lost peer traps instead of hanging the GPU), and reduces the
//                   W uncertified counts;
//   merge         : K6 (pack + select, bfk_select.cu) reads the W lists straight from the local region.
//
// Parity double-buffering is enough: a rank can finish search s only after every peer has pushed s, so
// nobody can be more than one search ahead of anybody else.  The kernels of different ranks run on
// different GPUs, so waiting on one another cannot deadlock a single GPU's scheduler.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bfk_common.cuh"

namespace bfk {

constexpr int XCHG_MAX_WORLD = 16;

struct Xchg {
    int device = 0, world = 1, rank = 0;
    size_t slot_bytes = 0;     // capacity of one rank's payload (multiple of 16)
    size_t parity_bytes = 0;   // W*slot_bytes + flags
    unsigned char* local = nullptr;                 // this rank's region (cudaMalloc)
    unsigned char* peer[XCHG_MAX_WORLD] = {nullptr};  // region of rank p as mapped here (peer[rank] == local)
    unsigned long long seq = 0;
    unsigned long long* d_maxflag = nullptr;        // device word: max uncertified count of the last exchange,
                                                    // or XCHG_TIMED_OUT when a peer's lists never arrived
    unsigned long long* h_maxflag = nullptr;        // pinned mirror
    int64_t block_rows = 0;    // > 0: local ids are block-cyclic (bfk_xchg_set_id_map) and made global while pushed
    long long timeout_cycles = 0;
};

constexpr unsigned long long XCHG_TIMED_OUT = ~0ull;

struct XchgPeers {
    unsigned char* p[XCHG_MAX_WORLD];
};

// payload layout inside a slot: [header: 2 words (nflag, nwords)] [D: ceil(nq*k/2) words] [I: nq*k words]
__global__ void __launch_bounds__(256) xchg_push_kernel(XchgPeers peers, int world, int rank, size_t parity_off,
                                                        size_t slot_bytes, size_t flags_off, const float* __restrict__ D,
                                                        const int64_t* __restrict__ I, int64_t nk,
                                                        const int* __restrict__ nflag, unsigned long long seq,
                                                        int64_t block_rows) {
    const int64_t wd = (nk + 1) / 2;
    const int64_t nwords = 2 + wd + nk;
    const unsigned long long* Dw = reinterpret_cast<const unsigned long long*>(D);  // read pairs of floats
    for (int p = blockIdx.y; p < world; p += gridDim.y) {
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(peers.p[p] + parity_off + (size_t)rank * slot_bytes);
        for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += (int64_t)gridDim.x * blockDim.x) {
            unsigned long long v;
            if (w == 0) v = nflag ? (unsigned long long)*nflag : 0ull;
            else if (w == 1) v = (unsigned long long)nwords;
            else if (w < 2 + wd) {
                const int64_t j = w - 2;
                if (2 * j + 1 < nk) v = Dw[j];
                else v = (unsigned long long)__float_as_uint(D[2 * j]);  // odd tail
            } else {
                int64_t id = I[w - 2 - wd];
                // block-cyclic placement: local row l of rank r is global row ((l / B) * W + r) * B + l % B
                if (block_rows > 0 && id >= 0) id = ((id / block_rows) * world + rank) * block_rows + id % block_rows;
                v = (unsigned long long)id;
            }
            dst[w] = v;
        }
    }
    // every CTA: make its stores visible system-wide, then the last CTA to finish publishes the flags
    __threadfence_system();
    __shared__ unsigned int last;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int* counter = reinterpret_cast<unsigned int*>(peers.p[rank] + parity_off + flags_off + world * 8);
        const unsigned int total = gridDim.x * gridDim.y;
        last = (atomicAdd(counter, 1u) == total - 1) ? 1u : 0u;
        if (last) *counter = 0u;
    }
    __syncthreads();
    if (last && threadIdx.x < world) {
        __threadfence_system();
        unsigned long long* flag =
            reinterpret_cast<unsigned long long*>(peers.p[threadIdx.x] + parity_off + flags_off) + rank;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(seq) : "memory");
    }
}

// A peer that does not show up within the time limit (it died, or its process is stuck) must not take this
// GPU's context with it: the kernel gives up, reports XCHG_TIMED_OUT through the status word the host reads
// anyway, and the host raises.  (It used to trap, which made the rank lose its whole shard.)
__global__ void xchg_wait_kernel(unsigned char* local, int world, size_t parity_off, size_t slot_bytes, size_t flags_off,
                                 unsigned long long seq, unsigned long long* maxflag, long long timeout_cycles) {
    const int lane = threadIdx.x;
    unsigned long long nf = 0;
    if (lane < world) {
        const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(local + parity_off + flags_off) + lane;
        const long long t0 = clock64();
        unsigned long long v;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
            if (v >= seq) break;
            if (clock64() - t0 > timeout_cycles) {
                printf("libbfk: exchange timed out waiting for rank %d (seq %llu, saw %llu)\n", lane, seq, v);
                nf = XCHG_TIMED_OUT;
                break;
            }
            __nanosleep(200);
        }
        if (nf != XCHG_TIMED_OUT)
            nf = *reinterpret_cast<const unsigned long long*>(local + parity_off + (size_t)lane * slot_bytes);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nf = max(nf, __shfl_xor_sync(0xffffffffu, nf, o));
    if (lane == 0) *maxflag = nf;
}

}  // namespace bfk

using namespace bfk;

extern "C" {

// allocates the exchange region and returns its CUDA IPC handle (64 bytes) for the peers
int bfk_xchg_create(int device, int world, int rank, int64_t max_nq_times_k, void** out, uint8_t* ipc_handle_out) {
    if (!out || !ipc_handle_out) return fail("xchg_create: null argument");
    if (world < 1 || world > XCHG_MAX_WORLD || rank < 0 || rank >= world) return fail("xchg_create: bad world/rank");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail("no CUDA device available: libbfk has no CPU fallback");
    }
    DeviceGuard g(device);
    Xchg* x = new Xchg();
    x->device = device;
    x->world = world;
    x->rank = rank;
    const int64_t nk = max_nq_times_k;
    x->slot_bytes = (size_t)round_up((2 + (nk + 1) / 2 + nk) * 8, 256);
    x->parity_bytes = (size_t)world * x->slot_bytes + (size_t)round_up(world * 8 + 8, 256);
    const size_t total = 2 * x->parity_bytes;
    cudaIpcMemHandle_t h;
    if (cudaMalloc((void**)&x->local, total) != cudaSuccess || cudaMemset(x->local, 0, total) != cudaSuccess ||
        cudaMalloc((void**)&x->d_maxflag, 8) != cudaSuccess ||
        cudaHostAlloc((void**)&x->h_maxflag, 8, cudaHostAllocDefault) != cudaSuccess ||
        cudaIpcGetMemHandle(&h, x->local) != cudaSuccess) {
        const char* e = cudaGetErrorString(cudaGetLastError());
        delete x;
        return fail(std::string("xchg_create: ") + e);
    }
    cudaDeviceSynchronize();
    // default limit 60 s (first-call module loads, a paused peer and collective regrowth all fit), BFK_XCHG_TIMEOUT_S overrides
    {
        const char* e = getenv("BFK_XCHG_TIMEOUT_S");
        const double secs = e ? std::max(0.001, atof(e)) : 60.0;
        x->timeout_cycles = (long long)(secs * 2.0e9);
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    memcpy(ipc_handle_out, &h, 64);
    x->peer[rank] = x->local;
    *out = x;
    return 0;
}

// maps every peer's region; all_handles = [world][64] as gathered from bfk_xchg_create on each rank
int bfk_xchg_connect(void* handle, const uint8_t* all_handles) {
    Xchg* x = (Xchg*)handle;
    if (!x || !all_handles) return fail("xchg_connect: null argument");
    DeviceGuard g(x->device);
    for (int p = 0; p < x->world; p++) {
        if (p == x->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, all_handles + (size_t)p * 64, 64);
        void* ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(std::string("xchg_connect: cannot map the region of rank ") + std::to_string(p) + ": " +
                        cudaGetErrorString(e));
        }
        x->peer[p] = (unsigned char*)ptr;
    }
    return 0;
}

int64_t bfk_xchg_capacity(void* handle) {
    Xchg* x = (Xchg*)handle;
    if (!x) return -1;
    return (int64_t)((x->slot_bytes / 8 - 2) * 2 / 3);  // nq*k such that 2 + ceil(nk/2) + nk words fit
}

// push this rank's lists to every rank, wait for everybody's, merge.  D_local/I_local [nq, k] device;
// nflag_dev: device int with this rank's count of not-yet-settled queries (may be NULL).
// After the stream has run, *max_nflag (host, pinned mirror copied on `stream`) > 0 means some rank will
// rewrite results and the exchange must be repeated by ALL ranks after they settle.
int bfk_xchg_merge(void* handle, const float* D_local, const int64_t* I_local, int64_t nq, int64_t k, int metric,
                   const int* nflag_dev, float* D_out, int64_t* I_out, void* stream) {
    Xchg* x = (Xchg*)handle;
    if (!x) return fail("xchg_merge: null handle");
    if (nq <= 0 || k <= 0 || k > BFK_MAX_K) return fail("xchg_merge: bad shape");
    const int64_t nk = nq * k;
    if ((size_t)((2 + (nk + 1) / 2 + nk) * 8) > x->slot_bytes) return fail("xchg_merge: lists larger than the exchange slots");
    DeviceGuard g(x->device);
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned long long seq = x->seq + 1;  // committed below, once both kernels are enqueued
    const size_t parity_off = (size_t)(seq & 1ull) * x->parity_bytes;
    const size_t flags_off = (size_t)x->world * x->slot_bytes;
    XchgPeers peers;
    for (int p = 0; p < XCHG_MAX_WORLD; p++) peers.p[p] = x->peer[p];
    const int64_t nwords = 2 + (nk + 1) / 2 + nk;
    dim3 grid((unsigned)std::min<int64_t>((nwords + 255) / 256, 32), (unsigned)x->world);
    xchg_push_kernel<<<grid, 256, 0, st>>>(peers, x->world, x->rank, parity_off, x->slot_bytes, flags_off, D_local, I_local,
                                           nk, nflag_dev, seq, x->block_rows);
    BFK_CUDA(cudaGetLastError());
    xchg_wait_kernel<<<1, 32, 0, st>>>(x->local, x->world, parity_off, x->slot_bytes, flags_off, seq, x->d_maxflag,
                                       x->timeout_cycles);
    BFK_CUDA(cudaGetLastError());
    x->seq = seq;
    // the W lists now sit in the local region: part p = slot p, D after the 2-word header, I after D
    const unsigned char* base = x->local + parity_off;
    const int64_t wd = (nk + 1) / 2;
    BFK_TRY(bfk_merge_topk_strided(reinterpret_cast<const float*>(base + 16), (int64_t)(x->slot_bytes / 4),
                                   reinterpret_cast<const int64_t*>(base + 16 + wd * 8), (int64_t)(x->slot_bytes / 8),
                                   x->world, nq, k, metric, D_out, I_out, x->device, stream));
    BFK_CUDA(cudaMemcpyAsync(x->h_maxflag, x->d_maxflag, 8, cudaMemcpyDeviceToHost, st));
    return 0;
}

// valid after the stream of the last bfk_xchg_merge has been synchronised; -2: a peer's lists never arrived
int64_t bfk_xchg_last_max_nflag(void* handle) {
    Xchg* x = (Xchg*)handle;
    if (!x) return -1;
    return *x->h_maxflag == XCHG_TIMED_OUT ? -2 : (int64_t)*x->h_maxflag;
}

int bfk_xchg_set_id_map(void* handle, int64_t block_rows) {
    Xchg* x = (Xchg*)handle;
    if (!x || block_rows < 0) return fail("xchg_set_id_map: bad argument");
    x->block_rows = block_rows;
    return 0;
}

int bfk_xchg_free(void* handle) {
    Xchg* x = (Xchg*)handle;
    if (!x) return 0;
    DeviceGuard g(x->device);
    cudaDeviceSynchronize();
    for (int p = 0; p < x->world; p++)
        if (p != x->rank && x->peer[p]) cudaIpcCloseMemHandle(x->peer[p]);
    if (x->local) cudaFree(x->local);
    if (x->d_maxflag) cudaFree(x->d_maxflag);
    if (x->h_maxflag) cudaFreeHost(x->h_maxflag);
    delete x;
    return 0;
}

}  // extern "C"

// bfk_search_noalloc: exact search for callers that are not allowed to allocate.
//
// Who this is for: code that runs where device memory cannot be requested any more --
// a serving process that hands every byte of HBM to its own pool at start-up and forbids
// cudaMalloc afterwards, a latency-critical request handler that must not take the
// driver's allocation lock, or a host thread that only holds pre-registered buffers.
// bfk_search grows a per-index workspace with cudaMalloc the first time a shape needs it;
// this variant never does.  Its only scratch -- the per-CTA partial top-k lists that the
// scan kernel (K3a) writes and the merge kernel (K4) reads -- lives in the fixed array
// below, which is part of the library image: every GPU that loads libbfk.so holds these
// 1.25 MiB until the process exits, whether or not this entry point is ever called.
//
// Limits (rejected with an error otherwise): nq <= 8, k <= 128, at most 160 scan CTAs
// (a B200 has 148 SMs), device pointers only.
//
// Sharing: there is ONE array per GPU.  Every call on that GPU uses it and finds the
// previous call's lists in it.  Nothing in it has to start at zero: the scan of a call
// rewrites every slot [query][cta][0..k) that the merge of the same call reads (CTAs with no
// rows publish empty lists), so no clearing pass is needed and none is done.  Two calls whose
// kernels overlap on the GPU (different streams, or different host threads on different
// streams) race on the array and return garbage; calls on one stream are safe back to back.
//
// Set-up: `g_scratch` below is a host-side stand-in for the device array, not an address a
// kernel can use.  The first call on each device resolves the real device address with
// cudaGetSymbolAddress (with that device current) and caches it; make that first call
// outside any stream capture.  The scan kernel exists in one variant per (queries rounded up
// to 1/2/4/8, metric, storage type); the first launch of a variant lets the driver load its
// code and sets its shared-memory attribute.  A caller that must not trigger any driver work
// while serving runs each variant it will use once at start-up.
#include <mutex>

#include "bfk_common.cuh"

namespace bfk {

constexpr int NOALLOC_MAX_NQ = 8;
constexpr int NOALLOC_MAX_K = 128;
constexpr int NOALLOC_MAX_PARTS = 160;
constexpr size_t NOALLOC_SLOTS = (size_t)NOALLOC_MAX_NQ * NOALLOC_MAX_PARTS * NOALLOC_MAX_K;  // 163840

// [nq][parts][k] composite keys (ordered distance << 32 | row id); 1.25 MiB per GPU
static __device__ __align__(256) unsigned long long g_scratch[NOALLOC_SLOTS];

constexpr int NOALLOC_MAX_DEVICES = 64;
static uint64_t* g_scratch_addr[NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_noalloc_mu;

// device address of g_scratch on the current device (resolved once per device)
static int scratch_for_device(int device, uint64_t** out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail("search_noalloc: device index out of range");
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(std::string("search_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        }
        g_scratch_addr[device] = static_cast<uint64_t*>(p);
    }
    *out = g_scratch_addr[device];
    return 0;
}

}  // namespace bfk

extern "C" int bfk_search_noalloc(const bfk_index* idx, const float* q_dev, int64_t nq, int64_t k, float* D_dev,
                                  int64_t* I_dev, void* stream) {
    using namespace bfk;
    if (!idx) return fail("null index");
    if (nq < 0) return fail("search_noalloc: negative query count");
    if (k <= 0) return fail("search_noalloc: k must be positive");
    if (nq > NOALLOC_MAX_NQ)
        return fail("search_noalloc: at most " + std::to_string(NOALLOC_MAX_NQ) + " queries per call (got " +
                    std::to_string(nq) + "); use bfk_search for larger batches");
    if (k > NOALLOC_MAX_K)
        return fail("search_noalloc: k is limited to " + std::to_string(NOALLOC_MAX_K) + " (got " + std::to_string(k) +
                    "); use bfk_search for larger k");
    if (nq == 0) return 0;
    if (!q_dev || !D_dev || !I_dev) return fail("search_noalloc: null buffer");
    DeviceGuard g(idx->device);
    if (!g.ok) return fail("cudaSetDevice failed");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    // the index may be growing in another thread (add regrows and frees the row storage): read its
    // pointers and row count under its own mutex, like every other search (a host mutex, no allocation)
    std::lock_guard<std::mutex> lk_idx(idx->mu);

    size_t bytes = 0;
    int nparts = 0;
    BFK_TRY(scan_exact_workspace(idx, (int)nq, (int)k, &bytes, &nparts));
    if (nparts > NOALLOC_MAX_PARTS || (size_t)nq * nparts * k > NOALLOC_SLOTS)
        return fail("search_noalloc: this GPU launches " + std::to_string(nparts) +
                    " scan CTAs, more than the fixed scratch holds (" + std::to_string(NOALLOC_MAX_PARTS) + ")");
    uint64_t* scratch = nullptr;
    BFK_TRY(scratch_for_device(idx->device, &scratch));

    cudaStream_t st = (cudaStream_t)stream;
    // the scan reads q with row stride d and zero-fills the padding columns itself
    BFK_TRY(wait_rows_ready(idx, st));  // cudaStreamWaitEvent on an existing event: no allocation
    // no selector: the bitmap of a bfk_search_sel running in another thread is not this call's business
    BFK_TRY(launch_scan_exact(idx, q_dev, idx->d, (int)nq, (int)k, scratch, nparts, st, nullptr));
    BFK_TRY(launch_select(scratch, (int64_t)nparts * k, nq, (int)k, idx->metric, 0, D_dev, I_dev, k, st));
    return 0;
}

// K4 `select_topk` / K6 `merge_parts`: per query, the k smallest of m composite keys,
// written best-first as the (D, I) pair faiss returns.
//
// Stands in for faiss's heap_reorder + result hand-off at the end of IndexFlat::search
// (reached from llm-qa/main.py:101,117) and, for the row-sharded index, for the merge
// that follows the all-gather of per-GPU lists (BASELINE config 5; no reference
// counterpart).  Missing results follow faiss: I = -1, D = +FLT_MAX (L2) / -FLT_MAX (IP).
#include <cfloat>

#include "bfk_common.cuh"

namespace bfk {

constexpr int SEL_THREADS = 512;
constexpr int SEL_N = 4096;  // sort window (composites), 32 KiB of shared memory

__device__ void block_bitonic_sort(uint64_t* a, int n, int tid, int nthr) {
    for (int k2 = 2; k2 <= n; k2 <<= 1) {
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < (n >> 1); i += nthr) {
                int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
                int hi = lo + j;
                bool asc = (lo & k2) == 0;
                uint64_t x = a[lo], y = a[hi];
                if ((x > y) == asc) {
                    a[lo] = y;
                    a[hi] = x;
                }
            }
            __syncthreads();
        }
    }
}

// cand [nq][m]; if `gather` != nullptr the low 32 bits of a composite index gather[q*... ]
// (see pack_parts_kernel) instead of being the row id itself.
// list_len > 0 promises that cand[q] is a concatenation of ascending lists of that length (the
// per-CTA lists of K3a, the per-shard lists of K6).  Then a cheap bound on the k-th smallest
// exists: split the lists into k groups, take each group's smallest head, and the largest of
// those k distinct elements is >= the k-th smallest overall -- only the few composites under
// that bound need sorting (a 64-element sort instead of a 2048-element one at nq = 1).
__global__ void __launch_bounds__(SEL_THREADS) select_topk_kernel(const uint64_t* __restrict__ cand, int64_t m, int k,
                                                                  int list_len, int metric, int64_t id_offset,
                                                                  const int64_t* __restrict__ gather, int64_t gather_q_stride,
                                                                  int64_t gather_p_stride, int gather_k,
                                                                  float* __restrict__ D, int64_t* __restrict__ I,
                                                                  int64_t out_ld, const SelectNotify nt) {
    __shared__ uint64_t buf[SEL_N];
    __shared__ unsigned long long gmin[128];
    __shared__ unsigned long long bound_s;
    __shared__ int cnt_s;
    const int tid = threadIdx.x;
    const int64_t q = blockIdx.x;
    const uint64_t* src = cand + q * m;
    bool done = false;

    const int64_t nlists = list_len > 0 ? m / list_len : 0;
    if (list_len > 0 && k <= 128 && nlists >= k) {
        if (tid < k) gmin[tid] = COMP_EMPTY;
        if (tid == 0) {
            bound_s = 0ull;
            cnt_s = 0;
        }
        __syncthreads();
        for (int64_t l = tid; l < nlists; l += SEL_THREADS) atomicMin(&gmin[l % k], (unsigned long long)src[l * list_len]);
        __syncthreads();
        if (tid < k) atomicMax(&bound_s, gmin[tid]);
        __syncthreads();
        const uint64_t bound = bound_s;
        for (int64_t i = tid; i < m; i += SEL_THREADS) {
            const uint64_t c = src[i];
            if (c <= bound && c != COMP_EMPTY) {
                const int slot = atomicAdd(&cnt_s, 1);
                if (slot < SEL_N) buf[slot] = c;
            }
        }
        __syncthreads();
        const int cnt = cnt_s;
        if (cnt <= SEL_N) {
            int n = 32;
            while (n < cnt || n < k) n <<= 1;
            for (int i = cnt + tid; i < n; i += SEL_THREADS) buf[i] = COMP_EMPTY;
            __syncthreads();
            block_bitonic_sort(buf, n, tid, SEL_THREADS);
            done = true;
        }
        // (more survivors than the window holds: only with massive ties; take the general path)
    }
    if (!done && m <= SEL_N) {
        int n = 32;
        while (n < m || n < k) n <<= 1;
        __syncthreads();
        for (int i = tid; i < n; i += SEL_THREADS) buf[i] = (i < m) ? src[i] : COMP_EMPTY;
        __syncthreads();
        block_bitonic_sort(buf, n, tid, SEL_THREADS);
    } else if (!done) {
        // keep the best kpad at the front, fold the rest through the window
        int kpad = 32;
        while (kpad < k) kpad <<= 1;
        const int step = SEL_N - kpad;
        __syncthreads();
        for (int i = tid; i < kpad; i += SEL_THREADS) buf[i] = COMP_EMPTY;
        for (int64_t base = 0; base < m; base += step) {
            for (int i = tid; i < step; i += SEL_THREADS) buf[kpad + i] = (base + i < m) ? src[base + i] : COMP_EMPTY;
            __syncthreads();
            block_bitonic_sort(buf, SEL_N, tid, SEL_THREADS);
        }
    }
    for (int i = tid; i < k; i += SEL_THREADS) {
        const uint64_t c = (i < SEL_N) ? buf[i] : COMP_EMPTY;
        float dv;
        int64_t iv;
        if (c == COMP_EMPTY) {
            dv = (metric == BFK_METRIC_L2) ? FLT_MAX : -FLT_MAX;
            iv = -1;
        } else {
            float key = ord2f((uint32_t)(c >> 32));
            dv = (metric == BFK_METRIC_L2) ? key : -key;
            uint32_t lo = (uint32_t)(c & 0xffffffffu);
            if (gather) {
                uint32_t part = lo / (uint32_t)gather_k, slot = lo - part * (uint32_t)gather_k;
                iv = gather[(int64_t)part * gather_p_stride + q * gather_q_stride + slot];
            } else {
                iv = (int64_t)lo + id_offset;
            }
        }
        D[q * out_ld + i] = dv;
        I[q * out_ld + i] = iv;
    }
    if (nt.h_seq != nullptr) {
        // (D, I) are pinned host memory and the host is polling: every query's rows must have reached it before
        // the last CTA publishes the sequence number (it also re-arms the ticket for the next search)
        __threadfence_system();
        __syncthreads();
        if (tid == 0 && atomicAdd(nt.ticket, 1) == (int)gridDim.x - 1) {
            *nt.ticket = 0;
            __threadfence();
            *reinterpret_cast<volatile int*>(nt.h_seq) = nt.seq;
            __threadfence_system();
        }
    }
}

int launch_select(const uint64_t* cand, int64_t m_per_query, int64_t nq, int k, int metric, int64_t id_offset,
                  float* D, int64_t* I, int64_t out_ld, cudaStream_t st, const SelectNotify* nt) {
    if (nq <= 0) return 0;
    // every caller hands over ascending lists of length k (K3a's per-CTA lists)
    select_topk_kernel<<<(unsigned)nq, SEL_THREADS, 0, st>>>(cand, m_per_query, k, k, metric, id_offset, nullptr, 0, 0,
                                                            1, D, I, out_ld, nt ? *nt : SelectNotify{nullptr, nullptr, 0});
    BFK_CUDA(cudaGetLastError());
    return 0;
}

// (D, I) lists [nparts][nq][k] -> composites [nq][nparts*k]; low word = part*k + slot so that
// equal distances resolve by shard then by the shard's own (distance, id) order.
__global__ void __launch_bounds__(256) pack_parts_kernel(const float* __restrict__ Dp, int64_t d_part_stride,
                                                         const int64_t* __restrict__ Ip, int64_t i_part_stride,
                                                         int nparts, int64_t nq, int k, int metric,
                                                         uint64_t* __restrict__ cand) {
    const int64_t total = (int64_t)nparts * nq * k;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t p = i / (nq * k);
        int64_t r = i - p * nq * k;
        int64_t q = r / k;
        int s = (int)(r - q * k);
        uint64_t c = COMP_EMPTY;
        if (Ip[p * i_part_stride + r] >= 0) {
            const float dv = Dp[p * d_part_stride + r];
            float key = (metric == BFK_METRIC_L2) ? dv : -dv;
            c = make_comp(key, (uint32_t)(p * k + s));
        }
        cand[q * ((int64_t)nparts * k) + p * k + s] = c;
    }
}

int launch_pack_parts(const float* Dparts, const int64_t* Iparts, int nparts, int64_t nq, int k, int metric,
                      uint64_t* cand, cudaStream_t st) {
    const int64_t total = (int64_t)nparts * nq * k;
    if (total <= 0) return 0;
    int64_t blocks = std::min<int64_t>((total + 255) / 256, 148 * 8);
    pack_parts_kernel<<<(unsigned)blocks, 256, 0, st>>>(Dparts, nq * k, Iparts, nq * k, nparts, nq, k, metric, cand);
    BFK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace bfk

// Merge step of the row-sharded search.  Scratch: nq*nparts*k composites, allocated per call
// with the stream-ordered allocator (the payload is tiny; BASELINE config 5).  Part p's lists start at
// Dparts + p*d_part_stride / Iparts + p*i_part_stride (elements), so the caller can hand over the
// all-gather buffer as it is, without repacking.
extern "C" int bfk_merge_topk_strided(const float* Dparts, int64_t d_part_stride, const int64_t* Iparts,
                                      int64_t i_part_stride, int nparts, int64_t nq, int64_t k, int metric, float* D,
                                      int64_t* I, int device, void* stream) {
    using namespace bfk;
    if (nparts <= 0 || nq < 0 || k <= 0 || k > BFK_MAX_K) return fail("merge_topk: bad arguments");
    if (nq == 0) return 0;
    if (!Dparts || !Iparts || !D || !I) return fail("merge_topk: null buffer");
    if ((int64_t)nparts * k >= (1ll << 32)) return fail("merge_topk: too many candidates per query");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail("no CUDA device available: libbfk has no CPU fallback");
    }
    DeviceGuard g(device);
    cudaStream_t st = (cudaStream_t)stream;
    // Keep freed scratch in the device's default pool across synchronisations: with the default release
    // threshold of 0 every stream sync hands the block back to the driver and the next merge pays a real
    // allocation (measured: 0.46 ms per exchange instead of 0.05 ms).
    static bool pool_kept[64] = {false};
    if (device >= 0 && device < 64 && !pool_kept[device]) {
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        cudaGetLastError();
        pool_kept[device] = true;
    }
    uint64_t* cand = nullptr;
    BFK_CUDA(cudaMallocAsync((void**)&cand, sizeof(uint64_t) * (size_t)nq * nparts * k, st));
    const int64_t total = (int64_t)nparts * nq * k;
    const int64_t blocks = std::min<int64_t>((total + 255) / 256, 148 * 8);
    pack_parts_kernel<<<(unsigned)blocks, 256, 0, st>>>(Dparts, d_part_stride, Iparts, i_part_stride, nparts, nq, (int)k,
                                                        metric, cand);
    int rc = cudaGetLastError() == cudaSuccess ? 0 : fail("merge_topk: launch failed");
    if (!rc) {
        select_topk_kernel<<<(unsigned)nq, SEL_THREADS, 0, st>>>(cand, (int64_t)nparts * k, (int)k, (int)k, metric, 0,
                                                                Iparts, k, i_part_stride, (int)k, D, I, k, SelectNotify{nullptr, nullptr, 0});
        if (cudaGetLastError() != cudaSuccess) rc = fail("merge_topk: launch failed");
    }
    cudaFreeAsync(cand, st);
    return rc;
}

extern "C" int bfk_merge_topk(const float* Dparts, const int64_t* Iparts, int nparts, int64_t nq, int64_t k,
                              int metric, float* D, int64_t* I, int device, void* stream) {
    return bfk_merge_topk_strided(Dparts, nq * k, Iparts, nq * k, nparts, nq, k, metric, D, I, device, stream);
}

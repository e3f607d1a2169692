// K7 `range_scan_kernel`: exact range search (all rows within a radius), the LangChain-side extra
// faiss exposes as index.range_search(x, radius) -> (lims, D, I).
//
// Stands in for faiss IndexFlat::range_search (exhaustive_L2sqr / exhaustive_inner_product with a
// RangeSearchResultHandler): L2 keeps squared distances < radius, inner product keeps scores > radius,
// both strict.  Arithmetic is the exact path's: direct sum of (x - y)^2 in fp32 (the nq < 20 form), on the
// stored rows (fp32 master rows, or the bf16-stored rows upcast).
//
// The result size is data dependent, so the scan runs twice over the corpus: pass 0 counts per query,
// the host turns counts into `lims`, pass 1 recomputes and writes (D, I) at per-query cursors.  Rows are
// independent; a warp owns a row, keeps up to RS_QB queries' partial sums in registers, and the corpus is
// read once per pass per block of RS_QB queries -- HBM-bound like K3a, but not tuned like it: this is a
// side path (LangChain's retriever never calls it; score_threshold filtering happens after search).
#include <vector>

#include "bfk_common.cuh"

namespace bfk {

constexpr int RS_QB = 8;        // queries per corpus pass
constexpr int RS_THREADS = 256;

struct RangeParams {
    const float* x32;
    const __nv_bfloat16* x16;
    int64_t n;
    int d, ld32, ld16;
    const float* q;  // [nqb][ldq] fp32, zero padded
    int ldq, nqb;
    int is_l2;
    float radius;
    const unsigned char* sel;
    unsigned long long* counts;   // [nqb] pass 0: totals; pass 1: cursors (start at lims)
    float* D;                     // pass 1 only
    int64_t* I;
    int64_t id_offset;
};

template <int PASS>
__global__ void __launch_bounds__(RS_THREADS) range_scan_kernel(const RangeParams p) {
    extern __shared__ float q_s[];  // [nqb][ldq]
    for (int i = threadIdx.x; i < p.nqb * p.ldq; i += RS_THREADS) q_s[i] = p.q[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * RS_THREADS + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * RS_THREADS) >> 5;
    for (int64_t row = warp0; row < p.n; row += nwarps) {
        if (p.sel && !((p.sel[row >> 3] >> (row & 7)) & 1u)) continue;
        float acc[RS_QB];
#pragma unroll
        for (int j = 0; j < RS_QB; j++) acc[j] = 0.f;
        for (int c = lane; c < p.d; c += 32) {
            const float x = p.x32 ? p.x32[(size_t)row * p.ld32 + c] : __bfloat162float(p.x16[(size_t)row * p.ld16 + c]);
#pragma unroll
            for (int j = 0; j < RS_QB; j++) {
                if (j < p.nqb) {
                    const float y = q_s[j * p.ldq + c];
                    if (p.is_l2) {
                        const float t = x - y;
                        acc[j] = fmaf(t, t, acc[j]);
                    } else {
                        acc[j] = fmaf(x, y, acc[j]);
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < RS_QB; j++) {
            if (j < p.nqb) {
                float v = acc[j];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                const bool keep = p.is_l2 ? (v < p.radius) : (v > p.radius);
                if (keep && lane == 0) {
                    const unsigned long long pos = atomicAdd(p.counts + j, 1ull);
                    if (PASS == 1) {
                        p.D[pos] = v;
                        p.I[pos] = row + p.id_offset;
                    }
                }
            }
        }
    }
}

struct RangeResult {
    int device;
    int64_t total;
    float* D;
    int64_t* I;
};

}  // namespace bfk

using namespace bfk;

extern "C" {

int bfk_range_search(bfk_index* idx, const float* q, int64_t nq, float radius, int q_is_device, void* stream,
                     int64_t* lims, void** result) {
    if (!idx) return fail("null index");
    if (nq < 0) return fail("range_search: negative query count");
    if (!lims || !result) return fail("range_search: null output");
    *result = nullptr;
    lims[0] = 0;
    if (nq > 0 && !q) return fail("range_search: null queries");
    DeviceGuard g(idx->device);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(idx->mu);
    BFK_TRY(search_tensor_finish(idx, st, nullptr));
    const int ldq = idx->ld32;
    float* q_dev = nullptr;
    unsigned long long* cnt_dev = nullptr;
    RangeResult* res = new RangeResult{idx->device, 0, nullptr, nullptr};
    std::vector<unsigned long long> counts((size_t)std::max<int64_t>(nq, 1), 0ull);
    int rc = 0;
    auto cleanup = [&]() {
        if (q_dev) cudaFree(q_dev);
        if (cnt_dev) cudaFree(cnt_dev);
    };
    auto bail = [&](const std::string& msg) {
        cleanup();
        if (res->D) cudaFree(res->D);
        if (res->I) cudaFree(res->I);
        delete res;
        return fail(msg);
    };
    if (nq > 0) {
        if (cudaMalloc((void**)&q_dev, sizeof(float) * (size_t)nq * ldq) != cudaSuccess ||
            cudaMalloc((void**)&cnt_dev, sizeof(unsigned long long) * (size_t)nq) != cudaSuccess)
            return bail("range_search: out of device memory");
        if (cudaMemsetAsync(q_dev, 0, sizeof(float) * (size_t)nq * ldq, st) != cudaSuccess ||
            cudaMemcpy2DAsync(q_dev, sizeof(float) * ldq, q, sizeof(float) * idx->d, sizeof(float) * idx->d, (size_t)nq,
                              q_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st) != cudaSuccess)
            return bail("range_search: query copy failed");
    }
    RangeParams p;
    p.x32 = idx->x32;
    p.x16 = idx->x16;
    p.n = idx->ntotal;
    p.d = idx->d;
    p.ld32 = idx->ld32;
    p.ld16 = idx->ld16;
    p.ldq = ldq;
    p.is_l2 = idx->metric == BFK_METRIC_L2 ? 1 : 0;
    p.radius = radius;
    p.sel = nullptr;
    p.D = nullptr;
    p.I = nullptr;
    p.id_offset = 0;
    const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>((idx->ntotal + 7) / 8, (int64_t)idx->num_sms * 8));
    for (int pass = 0; pass < 2 && !rc && nq > 0 && idx->ntotal > 0; pass++) {
        if (pass == 0) {
            if (cudaMemsetAsync(cnt_dev, 0, sizeof(unsigned long long) * (size_t)nq, st) != cudaSuccess)
                return bail("range_search: memset failed");
        } else {
            // cursors start at lims
            std::vector<unsigned long long> cur((size_t)nq);
            for (int64_t i = 0; i < nq; i++) cur[(size_t)i] = (unsigned long long)lims[i];
            if (cudaMemcpyAsync(cnt_dev, cur.data(), sizeof(unsigned long long) * (size_t)nq, cudaMemcpyHostToDevice, st) !=
                    cudaSuccess ||
                cudaStreamSynchronize(st) != cudaSuccess)
                return bail("range_search: cursor copy failed");
            if (res->total == 0) break;
            if (cudaMalloc((void**)&res->D, sizeof(float) * (size_t)res->total) != cudaSuccess ||
                cudaMalloc((void**)&res->I, sizeof(int64_t) * (size_t)res->total) != cudaSuccess)
                return bail("range_search: out of device memory for the result");
            p.D = res->D;
            p.I = res->I;
        }
        for (int64_t q0 = 0; q0 < nq; q0 += RS_QB) {
            p.q = q_dev + (size_t)q0 * ldq;
            p.nqb = (int)std::min<int64_t>(RS_QB, nq - q0);
            p.counts = cnt_dev + q0;
            const size_t smem = sizeof(float) * (size_t)p.nqb * ldq;
            if (pass == 0) range_scan_kernel<0><<<grid, RS_THREADS, smem, st>>>(p);
            else range_scan_kernel<1><<<grid, RS_THREADS, smem, st>>>(p);
            if (cudaGetLastError() != cudaSuccess) return bail("range_search: kernel launch failed");
        }
        if (pass == 0) {
            if (cudaMemcpyAsync(counts.data(), cnt_dev, sizeof(unsigned long long) * (size_t)nq, cudaMemcpyDeviceToHost, st) !=
                    cudaSuccess ||
                cudaStreamSynchronize(st) != cudaSuccess)
                return bail("range_search: kernel execution failed");
            for (int64_t i = 0; i < nq; i++) lims[i + 1] = lims[i] + (int64_t)counts[(size_t)i];
            res->total = lims[nq];
        }
    }
    if (nq > 0 && idx->ntotal == 0)
        for (int64_t i = 0; i < nq; i++) lims[i + 1] = 0;
    if (cudaStreamSynchronize(st) != cudaSuccess) return bail("range_search: kernel execution failed");
    cleanup();
    *result = res;
    return 0;
}

int64_t bfk_range_result_size(void* result) { return result ? ((RangeResult*)result)->total : 0; }

int bfk_range_result_copy(void* result, float* D, int64_t* I, int out_is_device) {
    if (!result) return fail("range_result_copy: null result");
    RangeResult* r = (RangeResult*)result;
    if (r->total == 0) return 0;
    if (!D || !I) return fail("range_result_copy: null output");
    DeviceGuard g(r->device);
    const cudaMemcpyKind kind = out_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    BFK_CUDA(cudaMemcpy(D, r->D, sizeof(float) * (size_t)r->total, kind));
    BFK_CUDA(cudaMemcpy(I, r->I, sizeof(int64_t) * (size_t)r->total, kind));
    return 0;
}

int bfk_range_result_free(void* result) {
    if (!result) return 0;
    RangeResult* r = (RangeResult*)result;
    DeviceGuard g(r->device);
    if (r->D) cudaFree(r->D);
    if (r->I) cudaFree(r->I);
    delete r;
    return 0;
}

}  // extern "C"

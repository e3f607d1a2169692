// K2 `append_cast_norm` and the small row utilities around it.
//
// K2 replaces IndexFlatCodes::add as reached from semantic-indexer/indexer.py:41
// (index.add(float32[n, d])): besides the fp32 master copy it produces the bf16 scan
// copy (zero padded to a multiple of 64 columns, the tcgen05 K-block) and the fp32
// squared norm of every row, and keeps max ||x||^2 for the rerank certificate.
// HBM-bound: reads 4*d bytes/row, writes 4*ld32 + 2*ld16 + 4 bytes/row.
#include "bfk_common.cuh"

namespace bfk {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// one warp per row; lanes stride the columns so every access is coalesced
template <bool STORE_F32>
__global__ void __launch_bounds__(256) append_rows_kernel(const float* __restrict__ src, int64_t src_ld, int64_t n,
                                                          int d, float* __restrict__ x32, int ld32,
                                                          __nv_bfloat16* __restrict__ x16, int ld16,
                                                          float* __restrict__ nrm, unsigned int* __restrict__ maxn) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    float local_max = 0.f, local_err = 0.f;
    // 16-byte loads / 16-byte fp32 stores / 8-byte bf16 stores, all loads of a row in flight before the first use
    // (d <= 512: one round; the scalar form below costs four times the instructions per byte and fell to 0.70 of
    // HBM bandwidth once the SM clock is power-capped under sustained load)
    const bool vec = (d & 3) == 0 && (src_ld & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0;
    for (int64_t r = warp; r < n; r += nwarps) {
        const float* s = src + r * src_ld;
        float ss = 0.f, se = 0.f;
        if (vec) {
            for (int c0 = 4 * lane; c0 < ld16; c0 += 512) {
                float4 v[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int c = c0 + 128 * u;
                    v[u] = c < d ? __ldg(reinterpret_cast<const float4*>(s + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int c = c0 + 128 * u;
                    if (c < ld16) {
                        const __nv_bfloat162 lo = __floats2bfloat162_rn(v[u].x, v[u].y), hi = __floats2bfloat162_rn(v[u].z, v[u].w);
                        const float2 flo = __bfloat1622float2(lo), fhi = __bfloat1622float2(hi);
                        if (STORE_F32) {
                            if (c < ld32) *reinterpret_cast<float4*>(x32 + r * ld32 + c) = v[u];
                            ss += v[u].x * v[u].x + v[u].y * v[u].y + v[u].z * v[u].z + v[u].w * v[u].w;
                            const float e0 = v[u].x - flo.x, e1 = v[u].y - flo.y, e2 = v[u].z - fhi.x, e3 = v[u].w - fhi.y;
                            se += e0 * e0 + e1 * e1 + e2 * e2 + e3 * e3;  // what the scan copy loses (exact in fp32)
                        } else {
                            ss += flo.x * flo.x + flo.y * flo.y + fhi.x * fhi.x + fhi.y * fhi.y;
                        }
                        uint2 packed;
                        packed.x = *reinterpret_cast<const unsigned int*>(&lo);
                        packed.y = *reinterpret_cast<const unsigned int*>(&hi);
                        *reinterpret_cast<uint2*>(x16 + r * ld16 + c) = packed;
                    }
                }
            }
        } else {
            for (int c = lane; c < ld16; c += 32) {
                float v = (c < d) ? __ldg(s + c) : 0.f;
                __nv_bfloat16 b = __float2bfloat16_rn(v);
                if (STORE_F32) {
                    if (c < ld32) x32[r * ld32 + c] = v;
                    ss += v * v;
                    const float e = v - __bfloat162float(b);
                    se += e * e;
                } else {
                    float vb = __bfloat162float(b);
                    ss += vb * vb;
                }
                x16[r * ld16 + c] = b;
            }
        }
        ss = warp_sum(ss);
        if (STORE_F32) se = warp_sum(se);
        if (lane == 0) nrm[r] = ss;
        local_max = fmaxf(local_max, ss);
        local_err = fmaxf(local_err, se);
    }
    // non-negative floats order like their bit patterns.  maxn[0] = max ||x||^2, maxn[1] = max ||x - bf16(x)||^2
    // (0 for a bf16-stored index, whose rows ARE the scan copy): both feed K5's error bound
    if (lane == 0 && local_max > 0.f) atomicMax(maxn, __float_as_uint(local_max));
    if (lane == 0 && local_err > 0.f) atomicMax(maxn + 1, __float_as_uint(local_err));
}

int launch_append_f32(bfk_index* idx, const float* src_dev, int64_t src_ld, int64_t n, cudaStream_t st) {
    if (n <= 0) return 0;
    const int warps_per_block = 8;
    int64_t blocks = std::min<int64_t>((n + warps_per_block - 1) / warps_per_block, (int64_t)idx->num_sms * 16);
    float* o32 = idx->x32 ? idx->x32 + (size_t)idx->ntotal * idx->ld32 : nullptr;
    __nv_bfloat16* o16 = idx->x16 + (size_t)idx->ntotal * idx->ld16;
    float* on = idx->nrm + idx->ntotal;
    if (idx->storage == BFK_STORE_F32)
        append_rows_kernel<true><<<(unsigned)blocks, 256, 0, st>>>(src_dev, src_ld, n, idx->d, o32, idx->ld32, o16,
                                                                   idx->ld16, on, idx->maxn);
    else
        append_rows_kernel<false><<<(unsigned)blocks, 256, 0, st>>>(src_dev, src_ld, n, idx->d, nullptr, 0, o16,
                                                                    idx->ld16, on, idx->maxn);
    BFK_CUDA(cudaGetLastError());
    return 0;
}

// bf16-stored rows -> packed fp32 [n, d] (reconstruct / write_index for BFK_STORE_BF16)
__global__ void __launch_bounds__(256) rows_bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ x16, int ld16,
                                                               int64_t n, int d, float* __restrict__ out) {
    const int64_t total = n * d;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / d;
        int c = (int)(i - r * d);
        out[i] = __bfloat162float(x16[r * ld16 + c]);
    }
}

int launch_rows_to_f32(const bfk_index* idx, int64_t i0, int64_t n, float* out_dev, cudaStream_t st) {
    if (n <= 0) return 0;
    int64_t total = n * idx->d;
    int64_t blocks = std::min<int64_t>((total + 255) / 256, (int64_t)idx->num_sms * 16);
    rows_bf16_to_f32_kernel<<<(unsigned)blocks, 256, 0, st>>>(idx->x16 + (size_t)i0 * idx->ld16, idx->ld16, n, idx->d,
                                                              out_dev);
    BFK_CUDA(cudaGetLastError());
    return 0;
}

// faiss.normalize_L2 (fvec_renorm_L2): x[i] *= 1/sqrt(||x[i]||^2) when the norm is > 0
__global__ void __launch_bounds__(256) renorm_rows_kernel(float* __restrict__ x, int64_t n, int d) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t r = warp; r < n; r += nwarps) {
        float* row = x + r * d;
        float ss = 0.f;
        for (int c = lane; c < d; c += 32) ss += row[c] * row[c];
        ss = warp_sum(ss);
        if (ss > 0.f) {
            float inv = 1.0f / sqrtf(ss);
            for (int c = lane; c < d; c += 32) row[c] *= inv;
        }
    }
}

// ---- id selector -> per-row bias of the tensor scan ------------------------------------------
// bias[i] = +inf for rows the selector rejects (their key can then never pass a threshold), else the
// metric's own bias (||x||^2 for L2, 0 for inner product).  One pass over n floats; the scan itself
// runs unchanged.
__global__ void __launch_bounds__(256) sel_bias_kernel(const unsigned char* __restrict__ sel, const float* __restrict__ nrm,
                                                       int64_t n, int64_t n_pad, int is_l2, float* __restrict__ bias) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * blockDim.x) {
        float b = INFINITY;
        if (i < n && ((sel[i >> 3] >> (i & 7)) & 1u)) b = is_l2 ? nrm[i] : 0.f;
        bias[i] = b;
    }
}

int launch_sel_bias(const unsigned char* sel, const float* nrm, int64_t n, int64_t n_pad, int is_l2, float* bias,
                    cudaStream_t st) {
    if (n_pad <= 0) return 0;
    const int64_t blocks = std::min<int64_t>((n_pad + 255) / 256, 148 * 16);
    sel_bias_kernel<<<(unsigned)blocks, 256, 0, st>>>(sel, nrm, n, n_pad, is_l2, bias);
    BFK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace bfk

extern "C" int bfk_normalize_l2(float* x, int64_t n, int d, int x_is_device, int device, void* stream) {
    using namespace bfk;
    if (n < 0 || d <= 0) return fail("normalize_L2: bad shape");
    if (n == 0) return 0;
    if (!x) return fail("normalize_L2: null input");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail("no CUDA device available: libbfk has no CPU fallback");
    }
    DeviceGuard g(device);
    cudaStream_t st = (cudaStream_t)stream;
    float* dx = x;
    if (!x_is_device) {
        BFK_CUDA(cudaMalloc((void**)&dx, sizeof(float) * (size_t)n * d));
        if (cudaMemcpyAsync(dx, x, sizeof(float) * (size_t)n * d, cudaMemcpyHostToDevice, st) != cudaSuccess) {
            cudaFree(dx);
            return fail("normalize_L2: H2D copy failed");
        }
    }
    int64_t blocks = std::min<int64_t>((n + 7) / 8, 148 * 16);
    renorm_rows_kernel<<<(unsigned)blocks, 256, 0, st>>>(dx, n, d);
    int rc = 0;
    if (cudaGetLastError() != cudaSuccess) rc = fail("normalize_L2: launch failed");
    if (!x_is_device) {
        if (!rc && (cudaMemcpyAsync(x, dx, sizeof(float) * (size_t)n * d, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                    cudaStreamSynchronize(st) != cudaSuccess))
            rc = fail("normalize_L2: D2H copy failed");
        cudaFree(dx);
    }
    return rc;
}

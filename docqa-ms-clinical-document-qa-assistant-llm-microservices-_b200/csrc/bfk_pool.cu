// K1 `pool_norm_cast`: masked mean-pool + L2-normalise (+ bf16 cast, norms, append).
//
// Stands in for the tail of SentenceTransformer.encode (models.Pooling(mean) followed by
// models.Normalize) as reached from semantic-indexer/indexer.py:37 and llm-qa/main.py:25:
//     e = sum_t tok[t]*m[t] / max(sum_t m[t], 1e-9);   v = e / max(||e||, 1e-12)
// and, when the output pointers are the tail of an index, for the index.add at
// indexer.py:41 fused behind it (no fp32 round trip through host memory).
//
// HBM-bound: T*H*sizeof(tok) + T*sizeof(mask) bytes in, H*4 (+ H*2 + 4) bytes out per
// chunk; masked-out tokens are not read at all.  One CTA per chunk: threads are laid
// out as (token group, 16-byte column vector), every token row is read with 128-bit
// coalesced loads, 8 loads in flight per thread.
#include "bfk_common.cuh"

namespace bfk {

template <typename T>
struct TokLoad;
template <>
struct TokLoad<float> {
    static constexpr int VEC = 4;
    __device__ static __forceinline__ void vec(const float* p, float (&v)[4]) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ static __forceinline__ float one(const float* p) { return __ldg(p); }
};
template <>
struct TokLoad<__nv_bfloat16> {
    static constexpr int VEC = 8;
    __device__ static __forceinline__ void vec(const __nv_bfloat16* p, float (&v)[8]) {
        uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ static __forceinline__ float one(const __nv_bfloat16* p) { return __bfloat162float(*p); }
};
template <>
struct TokLoad<__half> {
    static constexpr int VEC = 8;
    __device__ static __forceinline__ void vec(const __half* p, float (&v)[8]) {
        uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
            float2 f = __half22float2(h);
            v[2 * i] = f.x;
            v[2 * i + 1] = f.y;
        }
    }
    __device__ static __forceinline__ float one(const __half* p) { return __half2float(*p); }
};

__device__ __forceinline__ float load_mask(const void* mask, int mask_dtype, int64_t i) {
    switch (mask_dtype) {
        case BFK_DT_I64: return (float)reinterpret_cast<const long long*>(mask)[i];
        case BFK_DT_I32: return (float)reinterpret_cast<const int*>(mask)[i];
        default: return (float)reinterpret_cast<const unsigned char*>(mask)[i];
    }
}

__device__ __forceinline__ float block_sum(float v, float* red /*[32]*/, int tid, int nthr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    float s = 0.f;
    const int nw = (nthr + 31) >> 5;
    for (int w = 0; w < nw; w++) s += red[w];
    return s;
}

// VECTOR = true: H % VEC == 0 and tok is 16-byte aligned
template <typename TokT, bool VECTOR>
__global__ void __launch_bounds__(512) pool_norm_kernel(const TokT* __restrict__ tok, const void* __restrict__ mask,
                                                        int mask_dtype, int T, int H, int CV, int TG,
                                                        float* __restrict__ out32, int64_t out32_ld,
                                                        __nv_bfloat16* __restrict__ out16, int64_t out16_ld,
                                                        float* __restrict__ out_nrm, unsigned int* __restrict__ maxn) {
    constexpr int VEC = VECTOR ? TokLoad<TokT>::VEC : 1;
    extern __shared__ float sm[];
    float* mask_s = sm;                 // [T]
    float* red = mask_s + T;            // [TG][CV*VEC]
    float* scratch = red + TG * CV * VEC;  // [32]
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int64_t b = blockIdx.x;

    float msum = 0.f;
    for (int t = tid; t < T; t += nthr) {
        float m = load_mask(mask, mask_dtype, b * T + t);
        mask_s[t] = m;
        msum += m;
    }
    const float cnt = block_sum(msum, scratch, tid, nthr);  // also publishes mask_s

    const int cv = tid % CV, tg = tid / CV;
    float acc[VEC];
#pragma unroll
    for (int e = 0; e < VEC; e++) acc[e] = 0.f;
    if (tg < TG) {
        const TokT* base = tok + (size_t)b * T * H + (size_t)cv * VEC;
        constexpr int UN = 8;
        for (int t0 = tg; t0 < T; t0 += TG * UN) {
            float v[UN][VEC];
            float m[UN];
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int t = t0 + u * TG;
                m[u] = (t < T) ? mask_s[t] : 0.f;
                if (m[u] != 0.f) {
                    if constexpr (VECTOR) {
                        TokLoad<TokT>::vec(base + (size_t)t * H, v[u]);
                    } else {
                        v[u][0] = TokLoad<TokT>::one(base + (size_t)t * H);
                    }
                } else {
#pragma unroll
                    for (int e = 0; e < VEC; e++) v[u][e] = 0.f;
                }
            }
#pragma unroll
            for (int u = 0; u < UN; u++)
#pragma unroll
                for (int e = 0; e < VEC; e++) acc[e] = fmaf(v[u][e], m[u], acc[e]);
        }
#pragma unroll
        for (int e = 0; e < VEC; e++) red[(tg * CV + cv) * VEC + e] = acc[e];
    }
    __syncthreads();

    // columns: thread c < H folds the token groups
    for (int c = tid; c < H; c += nthr) {
        float s = 0.f;
        for (int g = 0; g < TG; g++) s += red[g * CV * VEC + c];
        red[c] = s / fmaxf(cnt, 1e-9f);  // group 0's slot now holds e[c]
    }
    __syncthreads();
    float ss = 0.f;
    for (int c = tid; c < H; c += nthr) ss += red[c] * red[c];
    ss = block_sum(ss, scratch, tid, nthr);
    const float inv = 1.0f / fmaxf(sqrtf(ss), 1e-12f);
    float ss2 = 0.f;
    const int64_t hi = out16 ? out16_ld : (out32 ? out32_ld : H);
    for (int64_t c = tid; c < hi; c += nthr) {
        float v = (c < H) ? red[c] * inv : 0.f;
        if (out32 && c < out32_ld) out32[b * out32_ld + c] = v;
        if (out16) {
            __nv_bfloat16 h = __float2bfloat16_rn(v);
            out16[b * out16_ld + c] = h;
            if (!out32) v = __bfloat162float(h);  // bf16-stored index: the norm of what is stored
        }
        ss2 += v * v;
    }
    if (out_nrm) {
        ss2 = block_sum(ss2, scratch, tid, nthr);
        if (tid == 0) {
            out_nrm[b] = ss2;
            if (maxn) atomicMax(maxn, __float_as_uint(ss2));
        }
    }
}

template <typename TokT>
static int launch_pool_t(const TokT* tok, const void* mask, int mask_dtype, int64_t B, int T, int H, float* out32,
                         int64_t out32_ld, __nv_bfloat16* out16, int64_t out16_ld, float* out_nrm, unsigned int* maxn,
                         cudaStream_t st) {
    constexpr int VEC = TokLoad<TokT>::VEC;
    const bool vec = (H % VEC == 0) && ((reinterpret_cast<uintptr_t>(tok) & 15u) == 0);
    const int vw = vec ? VEC : 1;
    const int CV = H / vw;
    if (CV > 512) return fail("pool_normalize: hidden size too large for this (unaligned) layout");
    int TG = std::max(1, std::min(std::min(16, 512 / CV), T));
    int threads = std::min(512, std::max(128, (int)round_up((int64_t)CV * TG, 32)));
    size_t smem = sizeof(float) * ((size_t)T + (size_t)TG * CV * vw + 32);
    if (smem > 200 * 1024) return fail("pool_normalize: sequence too long for shared memory");
    auto kern = vec ? pool_norm_kernel<TokT, true> : pool_norm_kernel<TokT, false>;
    if (smem > 48 * 1024) BFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)B, threads, smem, st>>>(tok, mask, mask_dtype, T, H, CV, TG, out32, out32_ld, out16, out16_ld,
                                            out_nrm, maxn);
    BFK_CUDA(cudaGetLastError());
    return 0;
}

int launch_pool(const void* tok, int tok_dtype, const void* mask, int mask_dtype, int64_t B, int T, int H,
                float* out32, int64_t out32_ld, __nv_bfloat16* out16, int64_t out16_ld, float* out_nrm,
                unsigned int* maxn, cudaStream_t st) {
    if (!tok || !mask) return fail("pool_normalize: null input");
    if (mask_dtype != BFK_DT_I64 && mask_dtype != BFK_DT_I32 && mask_dtype != BFK_DT_U8)
        return fail("pool_normalize: mask dtype must be int64, int32 or uint8/bool");
    if (B > 0x7fffffffll) return fail("pool_normalize: batch too large for one launch");
    switch (tok_dtype) {
        case BFK_DT_F32:
            return launch_pool_t((const float*)tok, mask, mask_dtype, B, T, H, out32, out32_ld, out16, out16_ld,
                                 out_nrm, maxn, st);
        case BFK_DT_BF16:
            return launch_pool_t((const __nv_bfloat16*)tok, mask, mask_dtype, B, T, H, out32, out32_ld, out16,
                                 out16_ld, out_nrm, maxn, st);
        case BFK_DT_F16:
            return launch_pool_t((const __half*)tok, mask, mask_dtype, B, T, H, out32, out32_ld, out16, out16_ld,
                                 out_nrm, maxn, st);
        default:
            return fail("pool_normalize: token dtype must be float32, float16 or bfloat16");
    }
}

}  // namespace bfk

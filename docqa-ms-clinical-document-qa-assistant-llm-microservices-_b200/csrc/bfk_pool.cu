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
#include <cstdlib>

#include "bfk_common.cuh"

namespace bfk {

// Token loads are split into "fetch the raw 16 bytes" and "convert + accumulate" so that the
// kernel can issue all of a batch's loads before touching any of them (8 x 16 B in flight per
// thread); fused load+convert made the compiler serialise them on one destination register.
// Eight predicated 128-bit read-only loads in ONE PTX block.  Written as separate statements, ptxas
// interleaves each load with the arithmetic on the previous one to save registers, which leaves a
// single 16-byte load in flight per thread and a quarter of HBM bandwidth; as one block all eight
// issue back to back.  A false predicate touches no memory and yields zeros.
__device__ __forceinline__ void ldg128_x8_if(const void* const (&p)[8], const unsigned (&on)[8], uint4 (&r)[8]) {
    asm volatile(
        "{\n\t.reg .pred q0, q1, q2, q3, q4, q5, q6, q7;\n\t"
        "setp.ne.u32 q0, %40, 0;\n\tsetp.ne.u32 q1, %41, 0;\n\tsetp.ne.u32 q2, %42, 0;\n\tsetp.ne.u32 q3, %43, 0;\n\t"
        "setp.ne.u32 q4, %44, 0;\n\tsetp.ne.u32 q5, %45, 0;\n\tsetp.ne.u32 q6, %46, 0;\n\tsetp.ne.u32 q7, %47, 0;\n\t"
        "mov.u32 %0, 0;\n\tmov.u32 %1, 0;\n\tmov.u32 %2, 0;\n\tmov.u32 %3, 0;\n\t"
        "mov.u32 %4, 0;\n\tmov.u32 %5, 0;\n\tmov.u32 %6, 0;\n\tmov.u32 %7, 0;\n\t"
        "mov.u32 %8, 0;\n\tmov.u32 %9, 0;\n\tmov.u32 %10, 0;\n\tmov.u32 %11, 0;\n\t"
        "mov.u32 %12, 0;\n\tmov.u32 %13, 0;\n\tmov.u32 %14, 0;\n\tmov.u32 %15, 0;\n\t"
        "mov.u32 %16, 0;\n\tmov.u32 %17, 0;\n\tmov.u32 %18, 0;\n\tmov.u32 %19, 0;\n\t"
        "mov.u32 %20, 0;\n\tmov.u32 %21, 0;\n\tmov.u32 %22, 0;\n\tmov.u32 %23, 0;\n\t"
        "mov.u32 %24, 0;\n\tmov.u32 %25, 0;\n\tmov.u32 %26, 0;\n\tmov.u32 %27, 0;\n\t"
        "mov.u32 %28, 0;\n\tmov.u32 %29, 0;\n\tmov.u32 %30, 0;\n\tmov.u32 %31, 0;\n\t"
        "@q0 ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%32];\n\t"
        "@q1 ld.global.nc.L1::no_allocate.v4.u32 {%4, %5, %6, %7}, [%33];\n\t"
        "@q2 ld.global.nc.L1::no_allocate.v4.u32 {%8, %9, %10, %11}, [%34];\n\t"
        "@q3 ld.global.nc.L1::no_allocate.v4.u32 {%12, %13, %14, %15}, [%35];\n\t"
        "@q4 ld.global.nc.L1::no_allocate.v4.u32 {%16, %17, %18, %19}, [%36];\n\t"
        "@q5 ld.global.nc.L1::no_allocate.v4.u32 {%20, %21, %22, %23}, [%37];\n\t"
        "@q6 ld.global.nc.L1::no_allocate.v4.u32 {%24, %25, %26, %27}, [%38];\n\t"
        "@q7 ld.global.nc.L1::no_allocate.v4.u32 {%28, %29, %30, %31}, [%39];\n\t}"
        : "=r"(r[0].x), "=r"(r[0].y), "=r"(r[0].z), "=r"(r[0].w), "=r"(r[1].x), "=r"(r[1].y), "=r"(r[1].z), "=r"(r[1].w),
          "=r"(r[2].x), "=r"(r[2].y), "=r"(r[2].z), "=r"(r[2].w), "=r"(r[3].x), "=r"(r[3].y), "=r"(r[3].z), "=r"(r[3].w),
          "=r"(r[4].x), "=r"(r[4].y), "=r"(r[4].z), "=r"(r[4].w), "=r"(r[5].x), "=r"(r[5].y), "=r"(r[5].z), "=r"(r[5].w),
          "=r"(r[6].x), "=r"(r[6].y), "=r"(r[6].z), "=r"(r[6].w), "=r"(r[7].x), "=r"(r[7].y), "=r"(r[7].z), "=r"(r[7].w)
        : "l"(p[0]), "l"(p[1]), "l"(p[2]), "l"(p[3]), "l"(p[4]), "l"(p[5]), "l"(p[6]), "l"(p[7]), "r"(on[0]), "r"(on[1]),
          "r"(on[2]), "r"(on[3]), "r"(on[4]), "r"(on[5]), "r"(on[6]), "r"(on[7]));
}

template <typename T>
struct TokLoad;
template <>
struct TokLoad<float> {
    static constexpr int VEC = 4;
    __device__ static __forceinline__ void fma(const uint4& t, float m, float (&acc)[4]) {
        acc[0] = fmaf(__uint_as_float(t.x), m, acc[0]);
        acc[1] = fmaf(__uint_as_float(t.y), m, acc[1]);
        acc[2] = fmaf(__uint_as_float(t.z), m, acc[2]);
        acc[3] = fmaf(__uint_as_float(t.w), m, acc[3]);
    }
    __device__ static __forceinline__ float one(const float* p) { return __ldg(p); }
};
template <>
struct TokLoad<__nv_bfloat16> {
    static constexpr int VEC = 8;
    __device__ static __forceinline__ void fma(const uint4& t, float m, float (&acc)[8]) {
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            acc[2 * i] = fmaf(__uint_as_float(w[i] << 16), m, acc[2 * i]);
            acc[2 * i + 1] = fmaf(__uint_as_float(w[i] & 0xffff0000u), m, acc[2 * i + 1]);
        }
    }
    __device__ static __forceinline__ float one(const __nv_bfloat16* p) { return __bfloat162float(*p); }
};
template <>
struct TokLoad<__half> {
    static constexpr int VEC = 8;
    __device__ static __forceinline__ void fma(const uint4& t, float m, float (&acc)[8]) {
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            acc[2 * i] = fmaf(f.x, m, acc[2 * i]);
            acc[2 * i + 1] = fmaf(f.y, m, acc[2 * i + 1]);
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
            float m[UN];
            if constexpr (VECTOR) {
                uint4 raw[UN];
                const void* ptr[UN];
                unsigned on[UN];
#pragma unroll
                for (int u = 0; u < UN; u++) {
                    const int t = t0 + u * TG;
                    m[u] = (t < T) ? mask_s[t] : 0.f;
                    on[u] = m[u] != 0.f;
                    ptr[u] = base + (size_t)min(t, T - 1) * H;
                }
                ldg128_x8_if(ptr, on, raw);
#pragma unroll
                for (int u = 0; u < UN; u++) TokLoad<TokT>::fma(raw[u], m[u], acc);
            } else {
                float v[UN];
#pragma unroll
                for (int u = 0; u < UN; u++) {
                    const int t = t0 + u * TG;
                    m[u] = (t < T) ? mask_s[t] : 0.f;
                    v[u] = (m[u] != 0.f) ? TokLoad<TokT>::one(base + (size_t)t * H) : 0.f;
                }
#pragma unroll
                for (int u = 0; u < UN; u++) acc[0] = fmaf(v[u], m[u], acc[0]);
            }
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
    float ss2 = 0.f, se2 = 0.f;
    const int64_t hi = out16 ? out16_ld : (out32 ? out32_ld : H);
    for (int64_t c = tid; c < hi; c += nthr) {
        float v = (c < H) ? red[c] * inv : 0.f;
        if (out32 && c < out32_ld) out32[b * out32_ld + c] = v;
        if (out16) {
            __nv_bfloat16 h = __float2bfloat16_rn(v);
            out16[b * out16_ld + c] = h;
            if (!out32) v = __bfloat162float(h);  // bf16-stored index: the norm of what is stored
            else {
                const float e = v - __bfloat162float(h);  // what the scan copy loses against the fp32 master row
                se2 += e * e;
            }
        }
        ss2 += v * v;
    }
    if (out_nrm) {
        ss2 = block_sum(ss2, scratch, tid, nthr);
        se2 = block_sum(se2, scratch, tid, nthr);
        if (tid == 0) {
            out_nrm[b] = ss2;
            if (maxn) {
                atomicMax(maxn, __float_as_uint(ss2));
                if (se2 > 0.f) atomicMax(maxn + 1, __float_as_uint(se2));  // see append_rows_kernel
            }
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
    // ~192-thread CTAs: every chunk has a prologue (mask) and an epilogue (fold, normalise, store) during which
    // it streams nothing, so many small CTAs per SM whose phases interleave beat few large ones.  Measured at
    // T=256, H=384 (fraction of HBM): bf16 0.67 / 0.75 / 0.82 / 0.70 and ragged fp32 0.76 / 0.84 / 0.84 / 0.74
    // for 512 / 256 / 192 / 128 threads (BFK_POOL_THREADS overrides, for tuning)
    static const int tgt_env = getenv("BFK_POOL_THREADS") ? atoi(getenv("BFK_POOL_THREADS")) : 0;
    const int tgt = tgt_env > 0 ? tgt_env : 192;
    int TG = std::max(1, std::min(std::min(16, tgt / CV), T));
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

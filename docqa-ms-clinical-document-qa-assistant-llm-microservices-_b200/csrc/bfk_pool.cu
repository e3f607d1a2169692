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
#include "bfk_ptx.cuh"

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

// ---- streaming form: persistent CTAs, token rows arrive by bulk copies --------------------------------------
// The kernel above is one CTA per chunk with register-staged 128-bit loads: every chunk pays a prologue (mask
// read, block sum) and an epilogue (fold, normalise, store) during which its CTA streams nothing -- 0.83 of the
// measured HBM bandwidth on full bf16 chunks, 0.66 on ragged ones, whose smaller payload hides the bubbles even
// less.  Here a CTA lives for the whole call.  Warp 8 is the producer: it reads the attention mask of the NEXT
// chunk while the current one streams, finds its last valid token, and copies the valid prefix of the chunk --
// token rows are contiguous, so 16 rows are ONE `cp.async.bulk` -- into a ring of shared-memory stages with
// mbarrier completion.  Warps 0-7 consume: thread = (token group, 16-byte column vector), conflict-free LDS.128,
// fp32 accumulation, tokens whose mask is 0 inside the prefix are skipped (never multiplied: a masked row may
// hold anything).  While the consumers fold and store one chunk, the producer is already filling the ring with
// the next one, so HBM never sees the epilogue.  Measured (bench.py --workload c4, fraction of the measured copy
// bandwidth; the copy moves half reads and half writes, a pure read stream can exceed it): bf16 full 0.83 -> 1.04,
// bf16 ragged 0.66 -> 0.83, fp32 full 0.92 -> 1.08, fp32 ragged 0.85 -> 0.99.
constexpr int PS_CWARPS = 8;
constexpr int PS_CTHREADS = PS_CWARPS * 32;
constexpr int PS_THREADS = PS_CTHREADS + 32;
constexpr int PS_MAX_STAGES = 8;
constexpr int PS_MAX_T = 512;  // mask values of one chunk live in the producer's registers (16 per lane)

__device__ __forceinline__ void ps_consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(PS_CTHREADS) : "memory"); }

__device__ __forceinline__ float ps_consumer_sum(float v, float* red, int tid) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    ps_consumer_sync();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    ps_consumer_sync();
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < PS_CWARPS; w++) s += red[w];
    return s;
}

struct PoolStreamParams {
    const void* tok;
    const void* mask;
    int mask_dtype;
    int64_t B;
    int T, H, CV, TG;
    int stage_tok, stages, stage_bytes, row_bytes;
    float* out32;
    int64_t out32_ld;
    __nv_bfloat16* out16;
    int64_t out16_ld;
    float* out_nrm;
    unsigned int* maxn;
};

template <typename TokT>
__global__ void __launch_bounds__(PS_THREADS, 2) pool_stream_kernel(const PoolStreamParams p) {
    constexpr int VEC = TokLoad<TokT>::VEC;
    extern __shared__ __align__(128) unsigned char ps_smem[];
    unsigned char* ring = ps_smem;
    float* mask_f = reinterpret_cast<float*>(ring + (size_t)p.stages * p.stage_bytes);  // [2][T]
    float* red = mask_f + 2 * p.T;                                                    // [TG][CV*VEC]
    float* scratch = red + p.TG * p.CV * VEC;                                         // [32]
    float* meta_cnt = scratch + 32;                                                   // [2] sum of the mask
    int* meta_tmax = reinterpret_cast<int*>(meta_cnt + 2);                            // [2] last valid token + 1
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(meta_tmax + 2);                  // (8-byte aligned by the host's sizes)
    uint64_t* empty_bar = full_bar + PS_MAX_STAGES;
    uint64_t* meta_ready = empty_bar + PS_MAX_STAGES;  // [2] the producer has published mask_f / meta of this parity
    uint64_t* mask_free = meta_ready + 2;              // [2] the consumers are done with them
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int i = 0; i < p.stages; i++) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], PS_CTHREADS);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&meta_ready[i], 1);
            mbar_init(&mask_free[i], PS_CTHREADS);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == PS_CWARPS) {
        // ===================== producer =====================
        float mreg[PS_MAX_T / 32];
        auto fetch_mask = [&](int64_t b) {
#pragma unroll
            for (int j = 0; j < PS_MAX_T / 32; j++) {
                const int t = lane + 32 * j;
                mreg[j] = t < p.T ? load_mask(p.mask, p.mask_dtype, b * p.T + t) : 0.f;
            }
        };
        if ((int64_t)blockIdx.x < p.B) fetch_mask(blockIdx.x);
        uint32_t st = 0, ph = 0;
        int it = 0;
        for (int64_t b = blockIdx.x; b < p.B; b += gridDim.x, it++) {
            const int par = it & 1;
            const uint32_t mph = (uint32_t)(it >> 1) & 1u;
            int tmax = 0;
            float cnt = 0.f;
#pragma unroll
            for (int j = 0; j < PS_MAX_T / 32; j++) {
                cnt += mreg[j];
                if (mreg[j] != 0.f) tmax = lane + 32 * j + 1;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
                tmax = max(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
            }
            mbar_wait(&mask_free[par], mph ^ 1u);  // the chunk that used this parity two chunks ago is finished
#pragma unroll
            for (int j = 0; j < PS_MAX_T / 32; j++) {
                const int t = lane + 32 * j;
                if (t < p.T) mask_f[par * p.T + t] = mreg[j];
            }
            if (lane == 0) {
                meta_cnt[par] = cnt;
                meta_tmax[par] = tmax;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&meta_ready[par]);
            const int64_t nb = b + gridDim.x;
            if (nb < p.B) fetch_mask(nb);  // in flight while this chunk's rows are being requested
            if (lane == 0) {
                const unsigned char* src = reinterpret_cast<const unsigned char*>(p.tok) + (size_t)b * p.T * p.row_bytes;
                for (int t0 = 0; t0 < tmax; t0 += p.stage_tok) {
                    const uint32_t bytes = (uint32_t)min(p.stage_tok, tmax - t0) * (uint32_t)p.row_bytes;
                    mbar_wait(&empty_bar[st], ph ^ 1u);
                    mbar_expect_tx(&full_bar[st], bytes);
                    bulk_g2s(ring + (size_t)st * p.stage_bytes, src + (size_t)t0 * p.row_bytes, bytes, &full_bar[st]);
                    if (++st == (uint32_t)p.stages) {
                        st = 0;
                        ph ^= 1u;
                    }
                }
            }
            __syncwarp();
        }
        return;
    }

    // ===================== consumers =====================
    const int cv = tid % p.CV, tg = tid / p.CV;
    const bool active = tg < p.TG;
    uint32_t st = 0, ph = 0;
    int it = 0;
    for (int64_t b = blockIdx.x; b < p.B; b += gridDim.x, it++) {
        const int par = it & 1;
        mbar_wait(&meta_ready[par], (uint32_t)(it >> 1) & 1u);
        const int tmax = meta_tmax[par];
        const float cnt = meta_cnt[par];
        const float* mf = mask_f + par * p.T;
        float acc[VEC];
#pragma unroll
        for (int e = 0; e < VEC; e++) acc[e] = 0.f;
        for (int t0 = 0; t0 < tmax; t0 += p.stage_tok) {
            const int rows = min(p.stage_tok, tmax - t0);
            mbar_wait(&full_bar[st], ph);
            if (active) {
                const unsigned char* base = ring + (size_t)st * p.stage_bytes + (size_t)cv * 16;
                // four rows per step: their LDS.128 are issued together, then folded in (a row whose mask is 0 is
                // neither read nor multiplied)
                for (int r0 = tg; r0 < rows; r0 += 4 * p.TG) {
                    uint4 v[4];
                    float m[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const int r = r0 + u * p.TG;
                        m[u] = r < rows ? mf[t0 + r] : 0.f;
                        if (m[u] != 0.f) v[u] = *reinterpret_cast<const uint4*>(base + (size_t)r * p.row_bytes);
                    }
#pragma unroll
                    for (int u = 0; u < 4; u++)
                        if (m[u] != 0.f) TokLoad<TokT>::fma(v[u], m[u], acc);
                }
            }
            mbar_arrive(&empty_bar[st]);
            if (++st == (uint32_t)p.stages) {
                st = 0;
                ph ^= 1u;
            }
        }
        // ---- fold the token groups, normalise, store (the producer is already streaming the next chunk) ----
        if (active) {
#pragma unroll
            for (int e = 0; e < VEC; e++) red[(tg * p.CV + cv) * VEC + e] = acc[e];
        }
        ps_consumer_sync();
        for (int c = tid; c < p.H; c += PS_CTHREADS) {
            float s = 0.f;
            for (int g = 0; g < p.TG; g++) s += red[g * p.CV * VEC + c];
            red[c] = s / fmaxf(cnt, 1e-9f);  // group 0's slot now holds e[c]
        }
        ps_consumer_sync();
        float ss = 0.f;
        for (int c = tid; c < p.H; c += PS_CTHREADS) ss += red[c] * red[c];
        ss = ps_consumer_sum(ss, scratch, tid);
        const float inv = 1.0f / fmaxf(sqrtf(ss), 1e-12f);
        float ss2 = 0.f, se2 = 0.f;
        const int64_t hi = p.out16 ? p.out16_ld : (p.out32 ? p.out32_ld : p.H);
        for (int64_t c = tid; c < hi; c += PS_CTHREADS) {
            float v = (c < p.H) ? red[c] * inv : 0.f;
            if (p.out32 && c < p.out32_ld) p.out32[b * p.out32_ld + c] = v;
            if (p.out16) {
                __nv_bfloat16 h = __float2bfloat16_rn(v);
                p.out16[b * p.out16_ld + c] = h;
                if (!p.out32) v = __bfloat162float(h);  // bf16-stored index: the norm of what is stored
                else {
                    const float e = v - __bfloat162float(h);  // what the scan copy loses against the fp32 master row
                    se2 += e * e;
                }
            }
            ss2 += v * v;
        }
        if (p.out_nrm) {
            ss2 = ps_consumer_sum(ss2, scratch, tid);
            se2 = ps_consumer_sum(se2, scratch, tid);
            if (tid == 0) {
                p.out_nrm[b] = ss2;
                if (p.maxn) {
                    atomicMax(p.maxn, __float_as_uint(ss2));
                    if (se2 > 0.f) atomicMax(p.maxn + 1, __float_as_uint(se2));
                }
            }
        }
        ps_consumer_sync();          // everyone is done with red / scratch / mask_f of this chunk
        mbar_arrive(&mask_free[par]);
    }
}

template <typename TokT>
static int launch_pool_t(const TokT* tok, const void* mask, int mask_dtype, int64_t B, int T, int H, float* out32,
                         int64_t out32_ld, __nv_bfloat16* out16, int64_t out16_ld, float* out_nrm, unsigned int* maxn,
                         cudaStream_t st) {
    constexpr int VEC = TokLoad<TokT>::VEC;
    const bool vec = (H % VEC == 0) && ((reinterpret_cast<uintptr_t>(tok) & 15u) == 0);
    // the streaming kernel: 16-byte aligned rows, the chunk's mask in the producer's registers, >= 2 ring stages
    static const bool stream_on = [] {
        const char* e = getenv("BFK_POOL_STREAM");
        return e ? atoi(e) != 0 : true;
    }();
    if (stream_on && vec && T <= PS_MAX_T && H / VEC <= PS_CTHREADS) {
        PoolStreamParams q;
        q.row_bytes = H * (int)sizeof(TokT);
        // stage = up to 32 token rows, at most 24 KB.  Measured at T=256, H=384 (fraction of the measured copy
        // bandwidth, bf16 full / bf16 ragged): 8 rows 0.60 / 0.49, 16 rows 0.97 / 0.76, 32 rows 1.04 / 0.83, 64 rows
        // 1.05 / 0.84 -- the per-stage barrier round trip has to be spread over enough rows per thread
        q.stage_tok = std::max(1, std::min(32, 24576 / q.row_bytes));
        q.stage_bytes = q.stage_tok * q.row_bytes;  // a multiple of 16
        q.CV = H / VEC;
        q.TG = std::max(1, std::min(PS_CTHREADS / q.CV, q.stage_tok));
        const size_t fixed = sizeof(float) * ((size_t)2 * T + (size_t)q.TG * H + 36) + 8 * (2 * PS_MAX_STAGES + 4) + 128;
        q.stages = (int)std::min<size_t>(PS_MAX_STAGES, (size_t)(98304) / (size_t)q.stage_bytes);
        if (q.stages >= 2) {
            const size_t smem = (size_t)q.stages * q.stage_bytes + fixed;
            q.tok = tok;
            q.mask = mask;
            q.mask_dtype = mask_dtype;
            q.B = B;
            q.T = T;
            q.H = H;
            q.out32 = out32;
            q.out32_ld = out32_ld;
            q.out16 = out16;
            q.out16_ld = out16_ld;
            q.out_nrm = out_nrm;
            q.maxn = maxn;
            auto kern = pool_stream_kernel<TokT>;
            static bool attr_set[64] = {false};  // per instantiation and device
            int dev = 0, sms = 148;
            BFK_CUDA(cudaGetDevice(&dev));
            if (dev < 0 || dev >= 64 || !attr_set[dev]) {
                BFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
                if (dev >= 0 && dev < 64) attr_set[dev] = true;
            }
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            if (smem <= 112 * 1024) {
                const unsigned grid = (unsigned)std::min<int64_t>(B, (int64_t)sms * 2);  // two resident CTAs per SM
                kern<<<grid, PS_THREADS, smem, st>>>(q);
                BFK_CUDA(cudaGetLastError());
                return 0;
            }
        }
    }
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

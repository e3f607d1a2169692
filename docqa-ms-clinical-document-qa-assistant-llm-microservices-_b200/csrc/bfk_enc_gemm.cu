// E2 `enc_gemm_kernel`: the linear layers of the sentence encoder on tcgen05.
//
// Stands in for the nn.Linear calls inside the 6 BertLayers that SentenceTransformer('all-MiniLM-L6-v2')
// runs for semantic-indexer/indexer.py:21,37 and llm-qa/main.py:25 (transformers BertSelfAttention
// query/key/value, BertSelfOutput.dense + LayerNorm, BertIntermediate.dense + GELU, BertOutput.dense +
// LayerNorm):
//     C[M, N] = A[M, K] . W[N, K]^T + bias      (A = token states, bf16 row-major; W = the layer's weight)
// with the op that follows fused into the epilogue:
//     EPI_BIAS    store bf16                                   (packed QKV projection, N = 1152)
//     EPI_GELU    exact (erf) GELU, store bf16                  (intermediate, N = 1536)
//     EPI_RES_LN  + residual row, LayerNorm(gamma, beta, eps)   (attention output / FFN output, N = 384)
//
// Tile = 128 rows x 384 columns: 384 is the model width, so a CTA owns complete rows and LayerNorm needs no
// cross-CTA (or even cross-thread) reduction -- epilogue thread r owns TMEM lane r = output row r, makes one
// pass over its 384 accumulator columns for mean / variance and a second one to normalise and store.  All N
// in the model are multiples of 384.  Per 64-wide K-block a stage holds the A block (16 KB) and three
// 128-row W blocks (48 KB), TMA-loaded with 128-byte swizzle; the MMA warp (converged, one elected lane)
// issues 4 x 3 tcgen05.mma (M=128, N=128, K=16) into three 128-column accumulators.  Persistent grid; the
// accumulator is single-buffered (384 of the 512 TMEM columns), so the MMAs of the next tile wait for the
// epilogue of this one while the TMA ring keeps prefetching -- the known cost of fusing the full-row
// LayerNorm.
#include <cuda.h>

#include "bfk_common.cuh"
#include "bfk_ptx.cuh"
#include "bfk_tmap.cuh"

namespace bfk {

constexpr int EG_BM = 128, EG_BN = 384, EG_BK = 64;
constexpr int EG_THREADS = 256;                       // 4 control warps + 4 epilogue warps
constexpr int EG_A_BYTES = EG_BM * EG_BK * 2;         // 16 KiB
constexpr int EG_B_BYTES = EG_BN * EG_BK * 2;         // 48 KiB (three 128-row blocks)
constexpr int EG_STAGE = EG_A_BYTES + EG_B_BYTES;     // 64 KiB
constexpr int EG_STAGES = 3;
constexpr int EG_SMEM = EG_STAGES * EG_STAGE + 1024 /*alignment*/ + 256 /*barriers*/;

enum { EPI_BIAS = 0, EPI_GELU = 1, EPI_RES_LN = 2 };

struct EncGemmParams {
    int M, N, K;
    const float* bias;               // [N]
    const __nv_bfloat16* residual;   // [M, 384] (EPI_RES_LN)
    const float* gamma;              // [384]
    const float* beta;               // [384]
    float eps;
    __nv_bfloat16* out;              // [M, ldo]
    int ldo;
};

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}

template <int EPI>
__global__ void __launch_bounds__(EG_THREADS, 1)
enc_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const EncGemmParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    unsigned char* smem = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    unsigned char* ring = smem;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + EG_STAGES * EG_STAGE);
    uint64_t* full_bar = bars;                    // [stages]
    uint64_t* empty_bar = bars + EG_STAGES;       // [stages]
    uint64_t* tfull_bar = bars + 2 * EG_STAGES;   // accumulator complete
    uint64_t* tempty_bar = tfull_bar + 1;         // accumulator drained (128 epilogue threads)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int KB = p.K / EG_BK;
    const int n_tiles = p.N / EG_BN;
    const int m_tiles = (p.M + EG_BM - 1) / EG_BM;
    const int tiles = m_tiles * n_tiles;

    if (tid == 0) {
        for (int i = 0; i < EG_STAGES; i++) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, 128);
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc<512>(tmem_slot);
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t st = 0, ph = 0;
            for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
                const int mt = tile / n_tiles, nt = tile - mt * n_tiles;
                for (int kb = 0; kb < KB; kb++) {
                    mbar_wait(&empty_bar[st], ph ^ 1u);
                    unsigned char* stage = ring + (size_t)st * EG_STAGE;
                    mbar_expect_tx(&full_bar[st], (uint32_t)EG_STAGE);
                    tma_load_2d(stage, &tmA, kb * EG_BK, mt * EG_BM, &full_bar[st]);
#pragma unroll
                    for (int nb = 0; nb < 3; nb++)
                        tma_load_2d(stage + EG_A_BYTES + nb * (128 * EG_BK * 2), &tmW, kb * EG_BK,
                                    nt * EG_BN + nb * 128, &full_bar[st]);
                    if (++st == EG_STAGES) {
                        st = 0;
                        ph ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (converged warp, one elected lane) =====================
        const uint32_t idesc = make_idesc_bf16(EG_BM, 128);
        const uint64_t desc_hi = make_smem_desc_sw128(0) & 0xffffffff00000000ull;
        const uint32_t desc_lo0 = (uint32_t)(make_smem_desc_sw128(0) & 0xffffffffull);
        const uint32_t ring_lo = desc_lo0 + ((smem_u32(ring) & 0x3ffffu) >> 4);
        uint32_t st = 0, ph = 0, tcount = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, tcount++) {
            mbar_wait(tempty_bar, (tcount & 1u) ^ 1u);  // the previous tile's epilogue has drained the accumulator
            tc_fence_after();
            for (int kb = 0; kb < KB; kb++) {
                mbar_wait(&full_bar[st], ph);
                tc_fence_after();
                const uint32_t a_lo = ring_lo + st * (EG_STAGE >> 4);
                const uint32_t b_lo = a_lo + (EG_A_BYTES >> 4);
                if (elect_one()) {
#pragma unroll
                    for (int kk = 0; kk < 4; kk++) {
                        const uint64_t da = desc_hi | (uint64_t)(a_lo + 2 * kk);
#pragma unroll
                        for (int nb = 0; nb < 3; nb++) {
                            const uint64_t db = desc_hi | (uint64_t)(b_lo + nb * ((128 * EG_BK * 2) >> 4) + 2 * kk);
                            umma_bf16(tmem_base + nb * 128, da, db, idesc, (kb | kk) != 0 ? 1u : 0u);
                        }
                    }
                    umma_commit(&empty_bar[st]);
                }
                __syncwarp();
                if (++st == EG_STAGES) {
                    st = 0;
                    ph ^= 1u;
                }
            }
            if (elect_one()) umma_commit(tfull_bar);
            __syncwarp();
        }
    } else if (warp >= 4) {
        // ===================== epilogue: thread == output row =====================
        const int row = tid - 128;  // TMEM lane
        const uint32_t lane_addr = tmem_base + ((uint32_t)((warp - 4) * 32) << 16);
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, tcount++) {
            const int mt = tile / n_tiles, nt = tile - mt * n_tiles;
            const int64_t m = (int64_t)mt * EG_BM + row;
            const bool live = m < p.M;
            const int n0 = nt * EG_BN;
            mbar_wait(tfull_bar, tcount & 1u);
            tc_fence_after();
            uint32_t r[32];
            if constexpr (EPI == EPI_RES_LN) {
                const uint4* res = reinterpret_cast<const uint4*>(p.residual + (size_t)(live ? m : 0) * EG_BN);
                float s1 = 0.f, s2 = 0.f;
#pragma unroll 1
                for (int c = 0; c < EG_BN; c += 32) {
                    tmem_ld_32x32(lane_addr + c, r);
                    uint4 rr[4];
#pragma unroll
                    for (int i = 0; i < 4; i++) rr[i] = res[(c >> 3) + i];
                    tmem_ld_wait();
                    const uint32_t* rw = reinterpret_cast<const uint32_t*>(rr);
#pragma unroll
                    for (int j = 0; j < 16; j++) {
                        const float v0 = __uint_as_float(r[2 * j]) + __ldg(p.bias + c + 2 * j) + __uint_as_float(rw[j] << 16);
                        const float v1 = __uint_as_float(r[2 * j + 1]) + __ldg(p.bias + c + 2 * j + 1) +
                                         __uint_as_float(rw[j] & 0xffff0000u);
                        s1 += v0 + v1;
                        s2 = fmaf(v0, v0, fmaf(v1, v1, s2));
                    }
                }
                const float mean = s1 * (1.0f / EG_BN);
                const float var = fmaxf(s2 * (1.0f / EG_BN) - mean * mean, 0.f);
                const float rstd = rsqrtf(var + p.eps);
                uint4* dst = reinterpret_cast<uint4*>(p.out + (size_t)(live ? m : 0) * p.ldo);
#pragma unroll 1
                for (int c = 0; c < EG_BN; c += 32) {
                    tmem_ld_32x32(lane_addr + c, r);
                    uint4 rr[4];
#pragma unroll
                    for (int i = 0; i < 4; i++) rr[i] = res[(c >> 3) + i];
                    tmem_ld_wait();
                    const uint32_t* rw = reinterpret_cast<const uint32_t*>(rr);
                    uint32_t o[16];
#pragma unroll
                    for (int j = 0; j < 16; j++) {
                        const float v0 = __uint_as_float(r[2 * j]) + __ldg(p.bias + c + 2 * j) + __uint_as_float(rw[j] << 16);
                        const float v1 = __uint_as_float(r[2 * j + 1]) + __ldg(p.bias + c + 2 * j + 1) +
                                         __uint_as_float(rw[j] & 0xffff0000u);
                        const float y0 = (v0 - mean) * rstd * __ldg(p.gamma + c + 2 * j) + __ldg(p.beta + c + 2 * j);
                        const float y1 = (v1 - mean) * rstd * __ldg(p.gamma + c + 2 * j + 1) + __ldg(p.beta + c + 2 * j + 1);
                        o[j] = pack_bf16x2(y0, y1);
                    }
                    if (live) {
#pragma unroll
                        for (int i = 0; i < 4; i++) dst[(c >> 3) + i] = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
                    }
                }
            } else {
                uint4* dst = reinterpret_cast<uint4*>(p.out + (size_t)(live ? m : 0) * p.ldo + n0);
#pragma unroll 1
                for (int c = 0; c < EG_BN; c += 32) {
                    tmem_ld_32x32(lane_addr + c, r);
                    tmem_ld_wait();
                    uint32_t o[16];
#pragma unroll
                    for (int j = 0; j < 16; j++) {
                        float v0 = __uint_as_float(r[2 * j]) + __ldg(p.bias + n0 + c + 2 * j);
                        float v1 = __uint_as_float(r[2 * j + 1]) + __ldg(p.bias + n0 + c + 2 * j + 1);
                        if constexpr (EPI == EPI_GELU) {
                            v0 = gelu_erf(v0);
                            v1 = gelu_erf(v1);
                        }
                        o[j] = pack_bf16x2(v0, v1);
                    }
                    if (live) {
#pragma unroll
                        for (int i = 0; i < 4; i++) dst[(c >> 3) + i] = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(tempty_bar);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

template <int EPI>
static int launch_enc_gemm_t(const CUtensorMap& tmA, const CUtensorMap& tmW, const EncGemmParams& p, int num_sms,
                             cudaStream_t st) {
    auto kern = enc_gemm_kernel<EPI>;
    static bool attr_set[64] = {false};
    int dev = 0;
    BFK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        BFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, EG_SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int tiles = ((p.M + EG_BM - 1) / EG_BM) * (p.N / EG_BN);
    const int grid = std::max(1, std::min(tiles, num_sms));
    kern<<<grid, EG_THREADS, EG_SMEM, st>>>(tmA, tmW, p);
    BFK_CUDA(cudaGetLastError());
    return 0;
}

// A [M, K] bf16 (ld = K), W [N, K] bf16 (ld = K); out [M, ldo] bf16
int launch_enc_gemm(int epi, const __nv_bfloat16* A, const __nv_bfloat16* W, int M, int N, int K, const float* bias,
                    const __nv_bfloat16* residual, const float* gamma, const float* beta, float eps,
                    __nv_bfloat16* out, int ldo, int num_sms, cudaStream_t st) {
    if (M <= 0) return 0;
    if (N % EG_BN != 0 || K % EG_BK != 0) return fail("encoder gemm: N must be a multiple of 384 and K of 64");
    if (epi == EPI_RES_LN && N != EG_BN) return fail("encoder gemm: the LayerNorm epilogue needs N == 384");
    CUtensorMap tmA, tmW;
    BFK_TRY(make_tmap(&tmA, A, M, K, EG_BM));
    BFK_TRY(make_tmap(&tmW, W, N, K, 128));
    EncGemmParams p;
    p.M = M;
    p.N = N;
    p.K = K;
    p.bias = bias;
    p.residual = residual;
    p.gamma = gamma;
    p.beta = beta;
    p.eps = eps;
    p.out = out;
    p.ldo = ldo;
    switch (epi) {
        case EPI_BIAS: return launch_enc_gemm_t<EPI_BIAS>(tmA, tmW, p, num_sms, st);
        case EPI_GELU: return launch_enc_gemm_t<EPI_GELU>(tmA, tmW, p, num_sms, st);
        case EPI_RES_LN: return launch_enc_gemm_t<EPI_RES_LN>(tmA, tmW, p, num_sms, st);
        default: return fail("encoder gemm: unknown epilogue");
    }
}

}  // namespace bfk

// E2 `enc_gemm_kernel`: the linear layers of the sentence encoder on tcgen05.
//
// Stands in for the nn.Linear calls inside the 6 BertLayers that SentenceTransformer('all-MiniLM-L6-v2')
// runs for semantic-indexer/indexer.py:21,37 and llm-qa/main.py:25 (transformers BertSelfAttention
// query/key/value, BertSelfOutput.dense + LayerNorm, BertIntermediate.dense + GELU, BertOutput.dense +
// LayerNorm):
//     C[M, N] = A[M, K] . W[N, K]^T + bias      (A = token states, bf16 row-major; W = the layer's weight)
// with the op that follows fused into the epilogue:
//     EPI_BIAS    store bf16                                   (packed QKV projection, N = 1152)
//     EPI_GELU    GELU (tanh form on MUFU.TANH; EPI_GELU_ERF: erf polynomial), store bf16   (intermediate, N = 1536)
//     EPI_RES_LN  + residual row, LayerNorm(gamma, beta, eps)   (attention output / FFN output, N = 384)
//
// Two tile shapes, one kernel (template<EPI, BN>):
//   BN = 384 (the model width) for the LayerNorm epilogue: a CTA owns complete rows, so LayerNorm needs no
//     cross-CTA reduction -- the two epilogue threads that share an output row (one per 192-column half)
//     exchange their partial sums through shared memory, pass 1 over the accumulator gives mean / variance and
//     writes v = acc + bias + residual back into TMEM, pass 2 normalises and stores.  384 of the 512 TMEM columns: single-buffered, the next tile's MMAs wait
//     for this tile's epilogue while the TMA ring keeps prefetching.
//   BN = 192 for the bias / GELU epilogues (all N in the model are multiples of 192): two accumulators, so
//     the MMAs of tile i+1 run under the epilogue of tile i.
// Per 64-wide K-block a stage holds the A block (16 KB) and the W block (BN rows), TMA-loaded with 128-byte
// swizzle; the MMA warp (converged, one elected lane) issues tcgen05.mma M=128, N=128 x3 or N=192, K=16.
// Persistent grid; 8 epilogue warps, thread = (output row = TMEM lane, column half), TMEM loads of the next
// 32 columns in flight while the current ones are finished.
//
// Round 2, in the order the measurements asked for them (BFK_ENC_GEMM_DEBUG prints where each role waits):
//   PAIR  two CTAs as one tcgen05.mma.cta_group::2 (M = 256): each CTA stages its own 128 rows of A and half of
//         every W block.  A lone CTA pulled 40 KB through L2 per 128 x 192 x 64 of MMA work and all four layer
//         shapes sat at the L2 slices' throughput limit (2.0 GB in 169 us for the intermediate layer).
//   ARES  (K = 384, 192-wide tiles; BFK_ENC_ARES=1, off by default: measured 5-7 % slower) the CTA's 128 x 384 slice
//         of A stays in shared memory while the pair sweeps the N tiles of its rows; the ring carries W blocks
//         only.  Each pair owns a contiguous range of the row-group-major tile sequence, so pairs change row groups
//         at different times.
//   All tiles store through TMA (one 32 x 32 box per chunk and warp) instead of 4 shared loads + 4 global
//         stores per lane: the LSU data pipe was the busiest unit of the QKV projection (67 % of peak), and pass 2
//         of the LayerNorm epilogue went from 7.4-8 k to 5.2-5.8 k cycles per tile.
//   Two-wide fp32 arithmetic (fma.rn.f32x2 = FFMA2, FADD2, FMUL2) in the bias / GELU / LayerNorm loops.
//   LayerNorm tiles pull their residual rows into L2 (and the first chunk into registers) while the tile's MMAs
//         run: their epilogue had half of its samples in long-scoreboard stalls on those loads.
#include <cuda.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "bfk_common.cuh"
#include "bfk_ptx.cuh"
#include "bfk_tmap.cuh"

namespace bfk {

constexpr int EG_BM = 128, EG_BK = 64;
constexpr int EG_THREADS = 384;                       // 4 control warps + 8 epilogue warps
constexpr int EG_EPI_THREADS = 256;
constexpr int EG_A_BYTES = EG_BM * EG_BK * 2;         // 16 KiB
constexpr int EG_BARS = 512;
constexpr int EG_ARES_KB = 6;                         // K-blocks of a resident A tile (K = 384)
constexpr int EG_FIXED = EG_BARS /*barriers*/ + 2 * 2 * 128 * 8 /*LayerNorm partial sums, double-buffered*/ +
                         3 * 384 * 4 /*bias, gamma, beta of a 384-wide tile*/;

template <int BN, bool PAIR, bool ARES = false>
struct EgShape {
    static constexpr int BNB = PAIR ? 192 : (BN == 384 ? 128 : BN);  // N of one tcgen05.mma
    static constexpr int NBLK = BN / BNB;
    static constexpr int NACC = BN == 192 ? 2 : 1;     // accumulators (256 TMEM columns apart)
    static constexpr int B_ROWS = PAIR ? BNB / 2 : BNB;  // rows of a W block this CTA stages (a pair splits them)
    static constexpr int B_BLOCK = B_ROWS * EG_BK * 2;
    static constexpr int B_BYTES = NBLK * B_BLOCK;
    // ARES: the CTA's 128 x 384 A tile stays in shared memory while the pair sweeps all N tiles of its rows; the
    // ring then carries W blocks only
    static constexpr int A_RES = ARES ? EG_ARES_KB * EG_A_BYTES : 0;
    static constexpr int STAGE = ARES ? B_BYTES : EG_A_BYTES + B_BYTES;
    static constexpr int STAGES = BN == 192 ? (ARES ? 7 : PAIR ? 6 : 4) : (PAIR ? 5 : 3);
    static constexpr int RING_END = A_RES + STAGES * STAGE;
    // per-warp staging between registers and global memory: 192-wide tiles store through TMA from two 2 KiB
    // buffers per warp, the LayerNorm tiles transpose through one
    static constexpr int STG_WARP = BN == 192 ? 4096 : 2048;
    static constexpr int SMEM = RING_END + 1024 /*alignment*/ + EG_FIXED + 8 * STG_WARP;
};

enum { EPI_BIAS = 0, EPI_GELU = 1, EPI_RES_LN = 2, EPI_GELU_ERF = 3 };

struct EncGemmParams {
    int M, N, K;
    const float* bias;               // [N]
    const __nv_bfloat16* residual;   // [M, 384] (EPI_RES_LN)
    const float* gamma;              // [384]
    const float* beta;               // [384]
    float eps;
    __nv_bfloat16* out;              // [M, ldo]
    int ldo;
    unsigned long long* dbg;         // BFK_ENC_GEMM_DEBUG: [grid][8] cycles each role spent waiting, or null
};

// exact-GELU through Abramowitz-Stegun 7.1.26 (|erf error| <= 1.5e-7, far below the bf16 rounding of the
// result): one reciprocal, one exp and a handful of FMAs instead of erff's ~25 instructions -- the GELU
// epilogue runs 128 x 192 of these per tile
__device__ __forceinline__ float gelu_erf(float x) {
    const float z = fabsf(x) * 0.70710678118654752f;
    const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));  // MUFU.RCP; __frcp_rn was 20 % of the epilogue's samples
    float poly = fmaf(1.061405429f, t, -1.453152027f);
    poly = fmaf(poly, t, 1.421413741f);
    poly = fmaf(poly, t, -0.284496736f);
    poly = fmaf(poly, t, 0.254829592f);
    const float erf_abs = 1.0f - poly * t * __expf(-z * z);
    return 0.5f * x * (1.0f + copysignf(erf_abs, x));
}

// GELU in its tanh form with the hardware tanh (MUFU.TANH, rel. error 2^-11): 5 instructions per element where
// the erf polynomial above needs ~19, and the GELU epilogue was issue-bound on exactly those (ncu: 43 % of the
// kernel's samples, stall_selected / not_selected).  |tanh form - erf form| <= 4.8e-4 absolute, attained near
// |x| = 2.6 where the bf16 result is rounded by up to 7.8e-3; below |x| = 0.5 the two differ by < 6e-5 against a
// rounding of >= 2.4e-4 -- the stored bf16 activation does not distinguish them.  BFK_ENC_GELU=erf selects the
// polynomial (tests run both against transformers' erf GELU).
__device__ __forceinline__ float gelu_tanh(float x) {
    const float u = x * fmaf(0.0356774081f, x * x, 0.7978845608f);
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
    const float hx = 0.5f * x;
    return fmaf(hx, t, hx);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}

#define EG_WAIT(bar, parity, slot)                                                  \
    do {                                                                            \
        const long long _t0 = p.dbg ? clock64() : 0;                                \
        mbar_wait(bar, parity);                                                     \
        if (p.dbg) dbg_acc[slot] += (unsigned long long)(clock64() - _t0);          \
    } while (0)

__device__ __forceinline__ void eg_epi_sync() { asm volatile("bar.sync 3, 256;" ::: "memory"); }

template <int EPI, int BN, bool PAIR, bool ARES>
__global__ void __launch_bounds__(EG_THREADS, 1)
enc_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                const __grid_constant__ CUtensorMap tmO, const EncGemmParams p) {
    static_assert(!ARES || (PAIR && BN == 192), "the resident-A form exists for paired 192-wide tiles");
    using S = EgShape<BN, PAIR, ARES>;
    constexpr uint32_t CS = PAIR ? 2u : 1u;  // CTAs per tile group
    constexpr int STAGES = S::STAGES, STAGE = S::STAGE, NACC = S::NACC;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    unsigned char* smem = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    unsigned char* a_res = smem;                  // ARES: [6 K-blocks][128 x 64] of this CTA's rows
    unsigned char* ring = smem + S::A_RES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::RING_END);
    uint64_t* full_bar = bars;                    // [stages]
    uint64_t* empty_bar = bars + STAGES;          // [stages]
    uint64_t* tfull_bar = bars + 2 * STAGES;      // [NACC] accumulator complete
    uint64_t* tempty_bar = tfull_bar + 2;         // [NACC] accumulator drained (256 epilogue threads)
    uint64_t* afull_bar = tempty_bar + 2;         // [6] ARES: K-block of the resident A tile has landed (both CTAs)
    uint64_t* afree_bar = afull_bar + EG_ARES_KB; // [6] ARES: the last MMA that reads it has completed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(afree_bar + EG_ARES_KB);
    float2* ln_part = reinterpret_cast<float2*>(smem + S::RING_END + EG_BARS);  // [2 parities][2 halves][128 rows]
    float* par_s = reinterpret_cast<float*>(smem + S::RING_END + EG_BARS + 4096);  // [3][384] bias | gamma | beta
    float* bias_s = reinterpret_cast<float*>(ln_part);  // bias / GELU kernels: the whole bias vector (N <= 2176 floats)
    unsigned char* stg_all = smem + S::RING_END + EG_BARS + 4096 + 3 * 384 * 4;    // [8 warps][32 rows x 64 B]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned long long dbg_acc[2] = {0, 0};
    const long long dbg_t0 = p.dbg ? clock64() : 0;
    const int KB = p.K / EG_BK;
    const int n_tiles = p.N / BN;
    // a pair works on 256 rows at a time: CTA rank r owns rows [256 g + 128 r, + 128) of group g
    const int m_tiles = (p.M + (int)CS * EG_BM - 1) / ((int)CS * EG_BM);
    const int tiles = m_tiles * n_tiles;
    const int first_tile = blockIdx.x / CS, tile_step = gridDim.x / CS;
    // the q-th tile of this CTA (pair): round-robin over all tiles, or -- ARES -- a contiguous range of the
    // row-group-major tile sequence, so that a pair sweeps the N tiles of a row group with its A tile in place.
    // The ranges start at different offsets inside a group (3072 tiles / 74 pairs), so the pairs change row
    // groups at different times: when they all did it together, 14 MB of A tiles were asked of HBM at once and
    // every group started with a 5.7 k-cycle wait (3.7 k remain: a block can only be requested once the previous
    // group's last MMA over it has completed, 2.3 k cycles before it is needed; pulling the next group's rows into
    // L2 a group ahead with prefetch.global.L2 from an idle warp changed nothing and was taken out again).  grp = row groups this pair has started before this tile's,
    // first / last = the tile is the first / last one of its row group inside the range.
    const int npairs = gridDim.x / CS;
    const int t_begin = (int)((int64_t)tiles * first_tile / npairs), t_end = (int)((int64_t)tiles * (first_tile + 1) / npairs);
    const int mg_begin = ARES ? t_begin / n_tiles : 0;
    int grp = 0;
    bool first = false, last = false;
    auto tile_at = [&](int q, int& mg, int& nt) {
        if constexpr (ARES) {
            const int t = t_begin + q;
            if (t >= t_end) return false;
            mg = t / n_tiles;
            nt = t - mg * n_tiles;
            grp = mg - mg_begin;
            first = q == 0 || nt == 0;
            last = t == t_end - 1 || nt == n_tiles - 1;
            return true;
        } else {
            const int tile = first_tile + q * tile_step;
            mg = tile / n_tiles;
            nt = tile - mg * n_tiles;
            return tile < tiles;
        }
    };

    if (tid == 0) {
        // pair: the leader's full / tempty barriers collect both CTAs; empty / tfull exist per CTA and are
        // signalled together by the leader's multicast commits
        for (int i = 0; i < STAGES; i++) {
            mbar_init(&full_bar[i], CS);
            mbar_init(&empty_bar[i], 1);
        }
        for (int a = 0; a < 2; a++) {
            mbar_init(&tfull_bar[a], 1);
            mbar_init(&tempty_bar[a], EG_EPI_THREADS * CS);
        }
        for (int i = 0; i < EG_ARES_KB; i++) {
            mbar_init(&afull_bar[i], CS);
            mbar_init(&afree_bar[i], 1);
        }
        fence_mbar_init();
    }
    if constexpr (EPI != EPI_RES_LN) {
        for (int i = tid; i < p.N; i += EG_THREADS) bias_s[i] = p.bias[i];  // read as shared-memory broadcasts
    }
    if constexpr (EPI == EPI_RES_LN) {
        // every tile of this kernel spans the same 384 columns: keep bias / gamma / beta on chip (read as
        // shared-memory broadcasts; as __ldg they queued behind the epilogue's other global traffic, stall_lg)
        for (int i = tid; i < BN; i += EG_THREADS) {
            par_s[i] = p.bias[i];
            par_s[BN + i] = p.gamma[i];
            par_s[2 * BN + i] = p.beta[i];
        }
    }
    if (warp == 2) {
        if constexpr (PAIR) tmem_alloc2<512>(tmem_slot);
        else tmem_alloc<512>(tmem_slot);
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        tma_prefetch_desc(&tmO);
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
    uint32_t crank = 0u;
    if constexpr (PAIR) crank = cluster_ctarank();
    const bool leader = crank == 0;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t st = 0, ph = 0;
            unsigned long long dbg_issue = 0;
            int mg, nt;
            for (int q = 0; tile_at(q, mg, nt); q++) {
                const int mt = mg * (int)CS + (int)crank;
                for (int kb = 0; kb < KB; kb++) {
                    if constexpr (ARES) {
                        if (first) {
                            // the next row group's K-block replaces this one as soon as the last N tile's MMAs over it
                            // have completed -- while the MMAs over the later K-blocks still run
                            const uint32_t gph = (uint32_t)grp & 1u;
                            EG_WAIT(&afree_bar[kb], gph ^ 1u, 1);
                            if (leader) mbar_expect_tx(&afull_bar[kb], 2u * (uint32_t)EG_A_BYTES);
                            tma_load_2d_2sm(a_res + kb * EG_A_BYTES, &tmA, kb * EG_BK, mt * EG_BM, &afull_bar[kb]);
                            if (!leader) mbar_arrive_remote(&afull_bar[kb], 0);
                        }
                    }
                    EG_WAIT(&empty_bar[st], ph ^ 1u, 0);
                    unsigned char* stage = ring + (size_t)st * STAGE;
                    if constexpr (ARES) {
                        const long long _t1 = p.dbg ? clock64() : 0;
                        if (leader) mbar_expect_tx(&full_bar[st], 2u * (uint32_t)STAGE);
                        tma_load_2d_2sm(stage, &tmW, kb * EG_BK, nt * BN + (int)crank * S::B_ROWS, &full_bar[st]);
                        if (!leader) mbar_arrive_remote(&full_bar[st], 0);
                        if (p.dbg) dbg_issue += (unsigned long long)(clock64() - _t1);
                    } else if constexpr (!PAIR) {
                        mbar_expect_tx(&full_bar[st], (uint32_t)STAGE);
                        tma_load_2d(stage, &tmA, kb * EG_BK, mt * EG_BM, &full_bar[st]);
#pragma unroll
                        for (int nb = 0; nb < S::NBLK; nb++)
                            tma_load_2d(stage + EG_A_BYTES + nb * S::B_BLOCK, &tmW, kb * EG_BK, nt * BN + nb * S::BNB,
                                        &full_bar[st]);
                    } else {
                        // each CTA stages its own 128 rows of A and its half of every W block in its own shared
                        // memory (the 256 x 192 MMA reads both halves); all bytes are accounted on the leader's barrier
                        if (leader) mbar_expect_tx(&full_bar[st], 2u * (uint32_t)STAGE);
                        tma_load_2d_2sm(stage, &tmA, kb * EG_BK, mt * EG_BM, &full_bar[st]);
#pragma unroll
                        for (int nb = 0; nb < S::NBLK; nb++)
                            tma_load_2d_2sm(stage + EG_A_BYTES + nb * S::B_BLOCK, &tmW, kb * EG_BK,
                                            nt * BN + nb * S::BNB + (int)crank * S::B_ROWS, &full_bar[st]);
                        if (!leader) mbar_arrive_remote(&full_bar[st], 0);
                    }
                    if (++st == STAGES) {
                        st = 0;
                        ph ^= 1u;
                    }
                }
            }
            if (p.dbg) {
                p.dbg[blockIdx.x * 8 + 0] = dbg_acc[0];  // producer: waiting for a free ring stage
                p.dbg[blockIdx.x * 8 + 1] = dbg_acc[1];  //           ... for a free resident-A block
                p.dbg[blockIdx.x * 8 + 5] = dbg_issue;   //           inside expect_tx + the W load (+ remote arrive)
            }
        }
    } else if (warp == 1 && leader) {
        // ===================== MMA issuer (converged warp, one elected lane; the pair's leader only) =====================
        const uint32_t idesc = make_idesc_bf16((int)CS * EG_BM, S::BNB);
        const uint64_t desc_hi = make_smem_desc_sw128(0) & 0xffffffff00000000ull;
        const uint32_t desc_lo0 = (uint32_t)(make_smem_desc_sw128(0) & 0xffffffffull);
        const uint32_t ring_lo = desc_lo0 + ((smem_u32(ring) & 0x3ffffu) >> 4);
        const uint32_t ares_lo = desc_lo0 + ((smem_u32(a_res) & 0x3ffffu) >> 4);
        unsigned long long dbg_afull = 0;
        uint32_t st = 0, ph = 0, tcount = 0;
        int mg, nt;
        for (int q = 0; tile_at(q, mg, nt); q++, tcount++) {
            const uint32_t a = NACC == 2 ? (tcount & 1u) : 0u;
            const uint32_t aph = NACC == 2 ? ((tcount >> 1) & 1u) : (tcount & 1u);
            EG_WAIT(&tempty_bar[a], aph ^ 1u, 0);  // the epilogue (of both CTAs) has drained this accumulator
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + a * 256;
            for (int kb = 0; kb < KB; kb++) {
                if constexpr (ARES) {
                    if (first) {
                        const long long _t2 = p.dbg ? clock64() : 0;
                        mbar_wait(&afull_bar[kb], (uint32_t)grp & 1u);
                        if (p.dbg) dbg_afull += (unsigned long long)(clock64() - _t2);
                    }
                }
                EG_WAIT(&full_bar[st], ph, 1);
                tc_fence_after();
                const uint32_t a_lo = ARES ? ares_lo + kb * (EG_A_BYTES >> 4) : ring_lo + st * (STAGE >> 4);
                const uint32_t b_lo = ARES ? ring_lo + st * (STAGE >> 4) : a_lo + (EG_A_BYTES >> 4);
                if (elect_one()) {
#pragma unroll
                    for (int kk = 0; kk < 4; kk++) {
                        const uint64_t da = desc_hi | (uint64_t)(a_lo + 2 * kk);
#pragma unroll
                        for (int nb = 0; nb < S::NBLK; nb++) {
                            const uint64_t db = desc_hi | (uint64_t)(b_lo + nb * (S::B_BLOCK >> 4) + 2 * kk);
                            if constexpr (PAIR) umma_bf16_2sm(d_tmem + nb * S::BNB, da, db, idesc, (kb | kk) != 0 ? 1u : 0u);
                            else umma_bf16(d_tmem + nb * S::BNB, da, db, idesc, (kb | kk) != 0 ? 1u : 0u);
                        }
                    }
                    if constexpr (PAIR) umma_commit_2sm(&empty_bar[st]);  // frees the stage in both CTAs
                    else umma_commit(&empty_bar[st]);
                    if constexpr (ARES) {
                        if (last) umma_commit_2sm(&afree_bar[kb]);  // this row group is done with the K-block
                    }
                }
                __syncwarp();
                if (++st == STAGES) {
                    st = 0;
                    ph ^= 1u;
                }
            }
            if (elect_one()) {
                if constexpr (PAIR) umma_commit_2sm(&tfull_bar[a]);  // accumulator complete, in both CTAs
                else umma_commit(&tfull_bar[a]);
            }
            __syncwarp();
        }
        if (p.dbg && lane == 0) {
            p.dbg[blockIdx.x * 8 + 2] = dbg_acc[0];  // MMA issuer: waiting for a drained accumulator
            p.dbg[blockIdx.x * 8 + 3] = dbg_acc[1];  //             ... for operands (ring stages)
            p.dbg[blockIdx.x * 8 + 4] = dbg_afull;   //             ... for resident-A blocks
            p.dbg[blockIdx.x * 8 + 6] = (unsigned long long)(clock64() - dbg_t0);
            p.dbg[blockIdx.x * 8 + 7] = tcount;
        }
    } else if (warp >= 4) {
        // ===================== epilogue: thread == (output row, column half) =====================
        constexpr int HC = BN / 2;     // columns per thread
        constexpr int NCH = HC / 32;   // 32-column chunks per thread
        const int et = tid - 128;
        const int row = et & 127, half = et >> 7;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(((warp - 4) & 3) * 32) << 16) + half * HC;
        // Stores go through a 2 KB per-warp transpose: thread = row would write 16 bytes to 32 different rows per
        // instruction (32 partial-sector requests; ncu: the epilogue sat in stall_lg with the LSU queue full).
        // Staged, four lanes cover 64 contiguous bytes of a row and one instruction touches 8 rows.
        unsigned char* stg = stg_all + (warp - 4) * S::STG_WARP;
        const int wrow0 = ((warp - 4) & 3) * 32;  // first tile row of this warp
        auto store32 = [&](const uint32_t (&o)[16], __nv_bfloat16* out_col0, int64_t m_tile0) {
            // o = this lane's row, 32 bf16 columns; out_col0 = out + first global column of the chunk
#pragma unroll
            for (int i = 0; i < 4; i++)
                *reinterpret_cast<uint4*>(stg + lane * 64 + ((i ^ ((lane >> 1) & 3)) << 4)) =
                    make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
            __syncwarp();
#pragma unroll
            for (int k8 = 0; k8 < 4; k8++) {
                const int rr = (lane >> 2) + 8 * k8, c = lane & 3;
                const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 64 + ((c ^ ((rr >> 1) & 3)) << 4));
                const int64_t mm = m_tile0 + wrow0 + rr;
                if (mm < p.M) *reinterpret_cast<uint4*>(out_col0 + (size_t)mm * p.ldo + c * 8) = v;
            }
            __syncwarp();
        };
        // 192-wide tiles: the same staged rows (the formula above is the 64-byte-swizzled [32 rows x 64 B] box) leave
        // through one TMA store per chunk instead of 4 shared loads + 4 global stores per lane: the epilogue's
        // shared-memory and store traffic had the LSU data pipe at 67 % of its peak (ncu, QKV projection) and set the
        // tile time, not the MMAs.  Two buffers per warp: a buffer is rewritten once the store before last has read it.
        uint32_t n_stores = 0;
        constexpr int NBUF = S::STG_WARP / 2048;  // staging buffers per warp
        auto store32_tma = [&](const uint32_t (&o)[16], int col0, int64_t m_tile0) {
            unsigned char* buf = stg + (n_stores % NBUF) * 2048;
            if (n_stores >= NBUF) {
                if (lane == 0) {
                    if constexpr (NBUF == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                }
                __syncwarp();
            }
#pragma unroll
            for (int i = 0; i < 4; i++)
                *reinterpret_cast<uint4*>(buf + lane * 64 + ((i ^ ((lane >> 1) & 3)) << 4)) =
                    make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&tmO, col0, (int)(m_tile0 + wrow0), buf);
                bulk_commit_group();
            }
            n_stores++;
        };
        // the mirror image for the residual rows: a coalesced read of 32 rows x 64 bytes (res_fetch, issued one
        // chunk ahead so its latency hides under the previous chunk's arithmetic), then every lane picks up its
        // own row through the staging buffer (res_stage).  Thread-per-row global reads were the LayerNorm
        // epilogue's long-scoreboard stalls.
        auto res_fetch = [&](const __nv_bfloat16* src_col0, int64_t m_tile0, uint4 (&g)[4]) {
#pragma unroll
            for (int k8 = 0; k8 < 4; k8++) {
                const int r8 = (lane >> 2) + 8 * k8, c = lane & 3;
                const int64_t mm = m_tile0 + wrow0 + r8;
                g[k8] = mm < p.M ? __ldg(reinterpret_cast<const uint4*>(src_col0 + (size_t)mm * BN + c * 8)) : make_uint4(0, 0, 0, 0);
            }
        };
        auto res_stage = [&](const uint4 (&g)[4], uint4 (&rr)[4]) {
#pragma unroll
            for (int k8 = 0; k8 < 4; k8++) {
                const int r8 = (lane >> 2) + 8 * k8, c = lane & 3;
                *reinterpret_cast<uint4*>(stg + r8 * 64 + ((c ^ ((r8 >> 1) & 3)) << 4)) = g[k8];
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; i++) rr[i] = *reinterpret_cast<const uint4*>(stg + lane * 64 + ((i ^ ((lane >> 1) & 3)) << 4));
            __syncwarp();
        };
        uint32_t tcount = 0;
        int mg, nt;
        for (int q = 0; tile_at(q, mg, nt); q++, tcount++) {
            const int mt = mg * (int)CS + (int)crank;
            const int n0 = nt * BN + half * HC;  // first global column of this thread
            const uint32_t a = NACC == 2 ? (tcount & 1u) : 0u;
            const uint32_t aph = NACC == 2 ? ((tcount >> 1) & 1u) : (tcount & 1u);
            uint4 gres[4];  // LayerNorm epilogue: residual chunk in flight
            if constexpr (EPI == EPI_RES_LN) {
                // This kernel has one accumulator, so these threads idle while the tile's MMAs run: the tile's residual
                // rows (128 x 768 bytes, last touched a kernel ago) are pulled into L2 now, and the first chunk into
                // registers -- the epilogue then never waits for HBM (ncu had it in long-scoreboard stalls for half
                // of its samples).
                const int64_t m0 = (int64_t)mt * EG_BM;
                for (int i = et; i < EG_BM * 6; i += EG_EPI_THREADS) {  // 6 lines of 128 bytes per row
                    const int64_t mm = m0 + i / 6;
                    if (mm < p.M)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(p.residual + (size_t)mm * BN + (i % 6) * 64));
                }
                res_fetch(p.residual + half * HC, m0, gres);
            }
            EG_WAIT(&tfull_bar[a], aph, 0);
            tc_fence_after();
            const uint32_t acc_addr = lane_addr + a * 256;
            uint32_t r[2][32];
            if constexpr (EPI == EPI_RES_LN) {
                if (lane == 0) bulk_wait_group_read0();  // the previous tile's stores have read the staging buffer
                __syncwarp();
                const __nv_bfloat16* res0 = p.residual + half * HC;  // first residual column of this thread
                const float* bias = par_s + half * HC;
                float s1 = 0.f, s2 = 0.f;
                uint64_t s1_2 = 0ull, s2_2 = 0ull;  // (even, odd) column partial sums
                const long long _e0 = p.dbg ? clock64() : 0;
                // one 32-column chunk of pass 1: v = acc + bias + residual, accumulate sum and sum of squares, and
                // write v back over the accumulator so that pass 2 needs neither the residual nor the bias again
                auto stats32 = [&](uint32_t (&rv)[32], int ch) {
                    uint4 rr[4];
                    res_stage(gres, rr);
                    if (ch + 1 < NCH) res_fetch(res0 + (ch + 1) * 32, (int64_t)mt * EG_BM, gres);
                    const uint32_t* rw = reinterpret_cast<const uint32_t*>(rr);
#pragma unroll
                    for (int j = 0; j < 16; j += 2) {  // four columns per round, two-wide fp32 arithmetic
                        const float4 bb = reinterpret_cast<const float4*>(bias + ch * 32)[j >> 1];
#pragma unroll
                        for (int k = 0; k < 2; k++) {
                            const uint64_t acc2 = f2_pack_bits(rv[2 * (j + k)], rv[2 * (j + k) + 1]);
                            const uint64_t b2 = k == 0 ? f2_pack(bb.x, bb.y) : f2_pack(bb.z, bb.w);
                            const uint64_t res2 = f2_pack_bits(rw[j + k] << 16, rw[j + k] & 0xffff0000u);
                            const uint64_t v2 = f2_add(f2_add(acc2, b2), res2);
                            s1_2 = f2_add(s1_2, v2);
                            s2_2 = f2_fma(v2, v2, s2_2);
                            f2_unpack_bits(v2, rv[2 * (j + k)], rv[2 * (j + k) + 1]);
                        }
                    }
                    tmem_st_32x32(acc_addr + ch * 32, rv);
                };
                // rolled over chunk pairs (two register sets, the next load in flight): unrolled over all six
                // chunks this epilogue spilled.  A register set is loaded again only after its store has drained.
                tmem_ld_32x32(acc_addr, r[0]);
#pragma unroll 1
                for (int ch = 0; ch < NCH; ch += 2) {
                    tmem_ld_wait();
                    tmem_st_wait();
                    tmem_ld_32x32(acc_addr + (ch + 1) * 32, r[1]);
                    stats32(r[0], ch);
                    tmem_ld_wait();
                    tmem_st_wait();
                    if (ch + 2 < NCH) tmem_ld_32x32(acc_addr + (ch + 2) * 32, r[0]);
                    stats32(r[1], ch + 1);
                }
                tmem_st_wait();
                {
                    float a0, a1, b0, b1;
                    f2_unpack(s1_2, a0, a1);
                    f2_unpack(s2_2, b0, b1);
                    s1 = a0 + a1;
                    s2 = b0 + b1;
                }
                const long long _e1 = p.dbg ? clock64() : 0;
                float2* part = ln_part + (tcount & 1u) * 256;
                part[half * 128 + row] = make_float2(s1, s2);
                eg_epi_sync();
                const long long _e2 = p.dbg ? clock64() : 0;
                const float2 other = part[(half ^ 1) * 128 + row];
                s1 += other.x;
                s2 += other.y;
                const float mean = s1 * (1.0f / BN);
                const float var = fmaxf(s2 * (1.0f / BN) - mean * mean, 0.f);
                const float rstd = rsqrtf(var + p.eps);
                const float* gamma = par_s + BN + half * HC;
                const float* beta = par_s + 2 * BN + half * HC;
                // (v - mean) rstd gamma + beta as two two-wide FMAs: v * rstd + (-mean rstd), then * gamma + beta
                const uint64_t a2 = f2_pack(rstd, rstd), b2 = f2_pack(-mean * rstd, -mean * rstd);
                auto norm32 = [&](const uint32_t (&rv)[32], int ch) {
                    uint32_t o[16];
#pragma unroll
                    for (int j = 0; j < 16; j += 2) {
                        const float4 gg = reinterpret_cast<const float4*>(gamma + ch * 32)[j >> 1];
                        const float4 be = reinterpret_cast<const float4*>(beta + ch * 32)[j >> 1];
#pragma unroll
                        for (int k = 0; k < 2; k++) {
                            const uint64_t t2 = f2_fma(f2_pack_bits(rv[2 * (j + k)], rv[2 * (j + k) + 1]), a2, b2);
                            const uint64_t y2 = f2_fma(t2, k == 0 ? f2_pack(gg.x, gg.y) : f2_pack(gg.z, gg.w),
                                                       k == 0 ? f2_pack(be.x, be.y) : f2_pack(be.z, be.w));
                            float y0, y1;
                            f2_unpack(y2, y0, y1);
                            o[j + k] = pack_bf16x2(y0, y1);
                        }
                    }
                    store32_tma(o, half * HC + ch * 32, (int64_t)mt * EG_BM);
                };
                tmem_ld_32x32(acc_addr, r[0]);
#pragma unroll 1
                for (int ch = 0; ch < NCH; ch += 2) {
                    tmem_ld_wait();
                    tmem_ld_32x32(acc_addr + (ch + 1) * 32, r[1]);
                    norm32(r[0], ch);
                    tmem_ld_wait();
                    if (ch + 2 < NCH) tmem_ld_32x32(acc_addr + (ch + 2) * 32, r[0]);
                    norm32(r[1], ch + 1);
                }
                if (p.dbg && et == 0) {  // row 0: cycles in pass 1, at the row-statistics barrier, in pass 2
                    unsigned long long* d = p.dbg + (size_t)gridDim.x * 8 + blockIdx.x * 8;
                    d[0] += (unsigned long long)(_e1 - _e0);
                    d[1] += (unsigned long long)(_e2 - _e1);
                    d[2] += (unsigned long long)(clock64() - _e2);
                }
            } else {
                const float* bias = bias_s + n0;
                tmem_ld_32x32(acc_addr, r[0]);
#pragma unroll
                for (int ch = 0; ch < NCH; ch++) {
                    tmem_ld_wait();
                    if (ch + 1 < NCH) tmem_ld_32x32(acc_addr + (ch + 1) * 32, r[(ch + 1) & 1]);
                    uint32_t o[16];
#pragma unroll
                    for (int j = 0; j < 16; j += 2) {  // four columns per round, two-wide fp32 arithmetic
                        const float4 bb = reinterpret_cast<const float4*>(bias + ch * 32)[j >> 1];
#pragma unroll
                        for (int k = 0; k < 2; k++) {
                            const uint64_t x2 = f2_add(f2_pack_bits(r[ch & 1][2 * (j + k)], r[ch & 1][2 * (j + k) + 1]),
                                                       k == 0 ? f2_pack(bb.x, bb.y) : f2_pack(bb.z, bb.w));
                            float v0, v1;
                            if constexpr (EPI == EPI_GELU) {
                                // gelu_tanh on the pair: u = x (c1 x^2 + c0), 0.5 x (1 + tanh u)
                                const uint64_t w2 = f2_fma(f2_mul(x2, x2), f2_pack(0.0356774081f, 0.0356774081f),
                                                           f2_pack(0.7978845608f, 0.7978845608f));
                                float u0, u1, t0, t1;
                                f2_unpack(f2_mul(x2, w2), u0, u1);
                                asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(u0));
                                asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(u1));
                                const uint64_t hx2 = f2_mul(x2, f2_pack(0.5f, 0.5f));
                                f2_unpack(f2_fma(hx2, f2_pack(t0, t1), hx2), v0, v1);
                            } else {
                                f2_unpack(x2, v0, v1);
                                if constexpr (EPI == EPI_GELU_ERF) {
                                    v0 = gelu_erf(v0);
                                    v1 = gelu_erf(v1);
                                }
                            }
                            o[j + k] = pack_bf16x2(v0, v1);
                        }
                    }
                    store32_tma(o, n0 + ch * 32, (int64_t)mt * EG_BM);
                }
            }
            tc_fence_before();
            if (leader) mbar_arrive(&tempty_bar[a]);
            else mbar_arrive_remote(&tempty_bar[a], 0);  // the MMA issuer lives in the leader CTA
        }
        if (lane == 0) bulk_wait_group_read0();  // the last stores have read their staging buffers

    }

    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // nobody leaves while the peer may still signal its barriers
    if (warp == 2) {
        tc_fence_after();
        if constexpr (PAIR) tmem_dealloc2<512>(tmem_base);
        else tmem_dealloc<512>(tmem_base);
    }
}

template <int EPI, int BN, bool PAIR, bool ARES>
static int launch_enc_gemm_t(const CUtensorMap& tmA, const CUtensorMap& tmW, const CUtensorMap& tmO, const EncGemmParams& p,
                             int num_sms, cudaStream_t st) {
    auto kern = enc_gemm_kernel<EPI, BN, PAIR, ARES>;
    using S = EgShape<BN, PAIR, ARES>;
    static bool attr_set[64] = {false};
    int dev = 0;
    BFK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        BFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    constexpr int CS = PAIR ? 2 : 1;
    const int m_groups = (p.M + CS * EG_BM - 1) / (CS * EG_BM);
    const int units = m_groups * (p.N / BN);  // tiles
    const int grid = CS * std::max(1, std::min(units, num_sms / CS));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(EG_THREADS);
    cfg.dynamicSmemBytes = S::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)CS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    static const bool debug = getenv("BFK_ENC_GEMM_DEBUG") != nullptr;
    if (debug) {  // synchronous: where each role of each CTA waited, on stderr
        EncGemmParams pd = p;
        BFK_CUDA(cudaMalloc((void**)&pd.dbg, (size_t)grid * 128));
        BFK_CUDA(cudaMemset(pd.dbg, 0, (size_t)grid * 128));
        BFK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmW, tmO, pd));
        BFK_CUDA(cudaStreamSynchronize(st));
        std::vector<unsigned long long> h((size_t)grid * 16);
        BFK_CUDA(cudaMemcpy(h.data(), pd.dbg, (size_t)grid * 128, cudaMemcpyDeviceToHost));
        if (EPI == EPI_RES_LN) {
            double e[3] = {0, 0, 0};
            for (int i = 0; i < grid; i += CS)
                for (int k = 0; k < 3; k++) e[k] += (double)h[(size_t)grid * 8 + (size_t)i * 8 + k];
            fprintf(stderr, "[bfk enc gemm] LayerNorm epilogue, row 0 of the leader CTAs, cycles per CTA: pass 1 %.0f, statistics barrier %.0f, pass 2 %.0f\n",
                    e[0] / (grid / CS), e[1] / (grid / CS), e[2] / (grid / CS));
        }
        double a[8] = {0};
        int nl = 0;
        for (int i = 0; i < grid; i += CS) {  // leaders
            for (int k = 0; k < 8; k++) a[k] += (double)h[(size_t)i * 8 + k];
            nl++;
        }
        fprintf(stderr,
                "[bfk enc gemm EPI=%d BN=%d pair=%d ares=%d M=%d N=%d K=%d] per leader CTA, cycles: life %.0f tiles %.1f | producer waits "
                "ring %.0f a-block %.0f, issuing W loads %.0f | mma waits accumulator %.0f ring stages %.0f a-blocks %.0f\n",
                EPI, BN, (int)PAIR, (int)ARES, p.M, p.N, p.K, a[6] / nl, a[7] / nl, a[0] / nl, a[1] / nl, a[5] / nl, a[2] / nl, a[3] / nl,
                a[4] / nl);
        return 0;
    }
    BFK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmW, tmO, p));
    BFK_CUDA(cudaGetLastError());
    return 0;
}

// mode: 0 = single CTAs, 1 = CTA pairs, 2 = CTA pairs with the A tile resident across the N tiles
template <int EPI, int BN>
static int launch_enc_gemm_p(int mode, const CUtensorMap& tmA, const CUtensorMap& tmW, const CUtensorMap& tmO,
                             const EncGemmParams& p, int num_sms, cudaStream_t st) {
    if constexpr (BN == 192) {
        if (mode == 2) return launch_enc_gemm_t<EPI, BN, true, true>(tmA, tmW, tmO, p, num_sms, st);
    }
    return mode >= 1 ? launch_enc_gemm_t<EPI, BN, true, false>(tmA, tmW, tmO, p, num_sms, st)
                     : launch_enc_gemm_t<EPI, BN, false, false>(tmA, tmW, tmO, p, num_sms, st);
}

// A [M, K] bf16 (ld = K), W [N, K] bf16 (ld = K); out [M, ldo] bf16
int launch_enc_gemm(int epi, const __nv_bfloat16* A, const __nv_bfloat16* W, int M, int N, int K, const float* bias,
                    const __nv_bfloat16* residual, const float* gamma, const float* beta, float eps,
                    __nv_bfloat16* out, int ldo, int num_sms, cudaStream_t st) {
    if (M <= 0) return 0;
    if (N % 192 != 0 || K % EG_BK != 0) return fail("encoder gemm: N must be a multiple of 192 and K of 64");
    if (N > 2176) return fail("encoder gemm: N > 2176 does not fit the on-chip bias vector");
    if (epi == EPI_RES_LN && N != 384) return fail("encoder gemm: the LayerNorm epilogue needs N == 384");
    // CTA pairs (cta_group::2, a 256-row tile group whose W blocks are split between the two CTAs) once there are
    // at least two 128-row tiles: a lone CTA re-reads 128 x 64 of A and BN x 64 of W per 128 x BN x 64 of MMA work
    // and all four layer shapes ran at the L2 slices' throughput limit, not the tensor pipe's.
    // BFK_ENC_PAIR=0 keeps single CTAs.
    static const bool pair_on = [] {
        const char* e = getenv("BFK_ENC_PAIR");
        return e ? atoi(e) != 0 : true;
    }();
    const bool pair = pair_on && M > EG_BM;
    // With K = 384 a CTA's whole 128 x K slice of A fits beside the ring: loaded once per row group instead of once
    // per N tile (QKV: 6 tiles, intermediate: 8), it halves what the pair pulls through L2.  Measured with everything
    // else in place it LOSES 5-7 % of CTA life (QKV 164.6 k vs 155.8 k cycles, intermediate 218.5 k vs 202.6 k:
    // profiles/r02_enc_ares_ab.log) -- what it saves in ring-stage waits it spends waiting for A blocks at every
    // row-group change -- so it is off unless BFK_ENC_ARES=1.
    static const bool ares_on = [] {
        const char* e = getenv("BFK_ENC_ARES");
        return e ? atoi(e) != 0 : false;
    }();
    const int mode = !pair ? 0 : (ares_on && K == EG_ARES_KB * EG_BK && epi != EPI_RES_LN && N > 192) ? 2 : 1;
    CUtensorMap tmA, tmW, tmO;
    BFK_TRY(make_tmap(&tmA, A, M, K, EG_BM));
    BFK_TRY(make_tmap_head(&tmO, out, M, ldo, 32));
    BFK_TRY(make_tmap(&tmW, W, N, K, pair ? 96 : (epi == EPI_RES_LN ? 128 : 192)));
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
    p.dbg = nullptr;
    switch (epi) {
        case EPI_BIAS: return launch_enc_gemm_p<EPI_BIAS, 192>(mode, tmA, tmW, tmO, p, num_sms, st);
        case EPI_GELU: {
            static const bool erf = [] {
                const char* e = getenv("BFK_ENC_GELU");
                return e && e[0] == 'e';
            }();
            return erf ? launch_enc_gemm_p<EPI_GELU_ERF, 192>(mode, tmA, tmW, tmO, p, num_sms, st)
                       : launch_enc_gemm_p<EPI_GELU, 192>(mode, tmA, tmW, tmO, p, num_sms, st);
        }
        case EPI_RES_LN: return launch_enc_gemm_p<EPI_RES_LN, 384>(mode, tmA, tmW, tmO, p, num_sms, st);
        default: return fail("encoder gemm: unknown epilogue");
    }
}

}  // namespace bfk

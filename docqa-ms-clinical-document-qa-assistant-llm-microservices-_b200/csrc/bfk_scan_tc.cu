// K3b `scan_tc_kernel` + K5 `rerank_kernel`: the tensor-core search path.
//
// Stands in for faiss IndexFlat::search's BLAS path (knn_L2sqr -> exhaustive_L2sqr_blas:
// sgemm + ||x||^2 + ||y||^2 - 2<x,y>, reached from llm-qa/main.py:101,117 when a batch of
// queries is searched) -- restructured for B200: the contraction runs on tcgen05 with a bf16
// scan copy of the corpus, selection is fused into the epilogue, and the survivors are
// re-scored exactly in fp32 with faiss's nq<20 arithmetic (direct sum((x-y)^2)).
//
// K3b (one CTA per SM, 12 warps, queries on M, corpus rows on N):
//   warp 0  TMA producer: the 128-query tile is loaded once (KB boxes of 64x128, SWIZZLE_128B)
//           and stays resident; corpus tiles stream as 64x256 boxes through an mbarrier ring
//   warp 1  one thread issues tcgen05.mma kind::f16 (bf16 x bf16 -> fp32), M=128 N=256 K=16,
//           4 per K-block; tcgen05.commit frees the smem stage / publishes the accumulator
//   warp 2  TMEM allocator (512 columns = two 128x256 fp32 accumulators, double buffered)
//   warps 4-11 epilogue, thread == (query row, 128-column half of the tile): tcgen05.ld 32x32b.x32,
//           key = ||x||^2 - 2*acc (L2) or -acc (IP) with the bias staged in smem, compare against
//           the thread's threshold (one warp vote per 4 columns skips the common no-survivor
//           case), append survivors to a per-thread pending list (coalesced [slot][thread] layout
//           in global scratch); when a list fills, the warp flushes: every thread bubbles its own
//           pending entries into a register-resident sorted top-KP, and the threshold tightens
//   CTA c works on query tile c % QT and corpus stream c / QT (a contiguous range of tiles), so
//   CTAs sharing a stream walk the same corpus rows together and share them through L2.
// Output per thread-stream (query, stream, column half): its KP best approximate keys + T, its final
//   threshold (every row it dropped has approximate key >= T).
//
// K5 (one CTA per query): merge the streams' lists by approximate key; A_k = k-th best;
//   everything with key <= A_k + 2*eps is re-scored exactly from the stored rows; eps bounds
//   |approximate - exact| from bf16 rounding: |<q,x> - <bf16 q, bf16 x>| <= (2^-8+) ||q|| ||x||.
//   The answer is certified exact when no dropped row can reach the cut (min_s T_s > A_k + 2*eps
//   and the re-score window was not exhausted); uncertified queries are flagged and re-run
//   through the exact scan K3a on the GPU -- never on the CPU.
//
// Roofline: HBM-bound while nq <~ 128 (2 bytes/element scanned, one pass per call); tensor-bound
// at large nq.  Algorithmic bytes per launch: N*ld16*2 + N*4 (norms, L2) + nq*d*2 + lists.
#include <cuda.h>

#include <type_traits>

#include <cfloat>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <utility>
#include <vector>

#include "bfk_common.cuh"
#include "bfk_ptx.cuh"
#include "bfk_tmap.cuh"

namespace bfk {

constexpr int TC_BM = 128;
constexpr int TC_BN = 256;
constexpr int TC_THREADS = 384;       // 4 control warps + 8 epilogue warps
constexpr int TC_EPI_THREADS = 256;
constexpr int TC_A_BLOCK = TC_BM * 128;  // bytes of one 64-column K-block of the query tile
constexpr int TC_B_BLOCK = TC_BN * 128;  // bytes of one 64-column K-block of a corpus tile
constexpr int TC_MAX_STAGES = 8;
constexpr int TC_BUFCAP = 16;            // pending slots per epilogue thread (shared memory)
constexpr int TC_SMEM_LIMIT = 232448;    // 227 KiB
constexpr int TC_RMAX = 512;             // exact re-scores per query at most
constexpr int TC_MERGE_N = 4096;
constexpr int TC_PEND_BYTES = TC_BUFCAP * 256 * 8;  // 32 KiB
constexpr int TC_TAIL_BYTES = 2 * TC_BN * 4 + (2 * TC_MAX_STAGES + 9) * 8 + 16 + 128 * 4;  // bias + barriers + tmem slot/flag + thr_sh

struct TcParams {
    const float* nrm;
    int64_t n;
    int KB;
    int QT;
    int S;
    int64_t tiles_total;
    int smem_bytes;      // dynamic shared memory given to the launch
    int is_l2;
    int nq;
    unsigned int* slots; // [nq_pad][NS] ordered key bits, slot (g*J + j) = min over the thread-streams of
                         // group g of their j-th best key: NS distinct rows, so max over the slots bounds
                         // the global NS-th (hence k-th, NS >= k) best key of the query from above
    int NS;              // slots per query (multiple of 4, >= k)
    int G;               // groups (power of two, G*J == NS, J <= KP)
    int J;               // ranks each thread publishes
    const float* qmargin;  // [nq_pad] 2*eps in key space (see prep_queries_kernel)
    uint64_t* out_cand;  // [QT*128][2*S][KP]  one list per thread-stream (query, stream, column half)
    float* out_thr;      // [QT*128][2*S]
    unsigned long long* dbg;  // optional [grid][8] cycle counters (BFK_TC_DEBUG=1), else nullptr
    int parked;          // producer/MMA waits use the suspend-time hint
    int a_resident;      // query tile resident in smem (else its K-blocks stream with the corpus)
    int use_bias;        // per-row bias is fetched (L2 norms, or the selector bias for either metric)
    int csize;           // 1: every CTA on its own; 2: CTA pairs (cta_group::2) -- two query tiles share one
                         // corpus stream, each CTA stages only half of every corpus tile
    unsigned long long* tstamp;  // optional [4] (globaltimer ns): [0] earliest scan CTA start, [1] latest scan CTA end,
                                 // [2] prep kernel start, [3] latest K5 CTA end; else nullptr
    int prepass;         // J == 1: tile 0 is first read for each thread's smallest key only (see the epilogue)
    uint64_t* pend_g;    // global pending lists [grid][TC_BUFCAP][256], or nullptr: lists live in shared memory.
                         // Long streams (several query tiles) append rarely and need the 32 KiB for a 4th
                         // corpus stage instead; short streams (nq <= 128) append often and keep them on chip.
};

// wait flavour for the single-thread roles + optional cycle accounting
// (DBG is a template parameter of the kernel: the shipped instantiation carries none of this)
#define TC_DBG (DBG && p.dbg != nullptr)
#define TC_WAIT(bar, parity, slot)                                   \
    do {                                                             \
        const long long _t0 = TC_DBG ? clock64() : 0;                \
        if (DBG && p.parked) mbar_wait_parked(bar, parity);          \
        else mbar_wait(bar, parity);                                 \
        if (TC_DBG) dbg_acc[slot] += (unsigned long long)(clock64() - _t0); \
    } while (0)

__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 2, 256;" ::: "memory"); }

// PAIR is a template parameter, not a run-time flag: a kernel that contains cta_group::2 instructions can
// only be launched with an even cluster size.
template <int KP, bool PAIR, bool DBG>
__global__ void __launch_bounds__(TC_THREADS, 1)
scan_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmX, const TcParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    const uint32_t pad = (1024u - (raw_addr & 1023u)) & 1023u;  // 0 when the base is already 1 KiB aligned
    unsigned char* smem = smem_raw + pad;
    // stages: whatever fits after the resident query tile and the fixed tail (host sizes for pad = 0;
    // a misaligned base just costs one stage)
    const int a_bytes = p.a_resident ? p.KB * TC_A_BLOCK : 0;
    constexpr bool pair = PAIR;
    const int b_bytes = pair ? TC_B_BLOCK / 2 : TC_B_BLOCK;  // a paired CTA stages 128 of the tile's 256 rows
    const int stage_bytes = p.a_resident ? b_bytes : b_bytes + TC_A_BLOCK;  // streamed: [B block | A block]
    const int pend_bytes = p.pend_g ? 0 : TC_PEND_BYTES;
    const int stages = min(TC_MAX_STAGES, (int)((p.smem_bytes - (int)pad - a_bytes - pend_bytes - TC_TAIL_BYTES) / stage_bytes));
    unsigned char* sA = smem;
    unsigned char* sB = sA + (size_t)a_bytes;
    uint64_t* pend_s = reinterpret_cast<uint64_t*>(sB + (size_t)stages * stage_bytes);  // [TC_BUFCAP][256]
    float* bias_s = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(pend_s) + pend_bytes);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(bias_s + 2 * TC_BN);
    uint64_t* empty_bar = full_bar + TC_MAX_STAGES;
    uint64_t* q_bar = empty_bar + TC_MAX_STAGES;
    uint64_t* tfull_bar = q_bar + 1;   // [2]
    uint64_t* tempty_bar = tfull_bar + 2;  // [2]
    uint64_t* nfull_bar = tempty_bar + 2;  // [2] norms of the tile have landed in bias slot a
    uint64_t* tlocal_bar = nfull_bar + 2;  // [2] THIS CTA's epilogue has released accumulator / bias slot a
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tlocal_bar + 2);
    volatile int* done_flag = reinterpret_cast<volatile int*>(tmem_slot + 1);
    float* thr_sh = reinterpret_cast<float*>(tmem_slot + 4);  // [128] bound on the global KP-th best key per query row

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long dbg_t0 = TC_DBG ? clock64() : 0;
    unsigned long long dbg_acc[4] = {0, 0, 0, 0};
    // CTA 0 also records a per-tile event timeline (first 64 tiles) behind the counters: [event][tile]
    unsigned long long* tl = (TC_DBG && blockIdx.x == 0) ? p.dbg + (size_t)gridDim.x * 8 : nullptr;
#define TC_MARK(ev, tile) do { if (tl && (tile) < 64) tl[(ev) * 64 + (tile)] = (unsigned long long)(clock64() - dbg_t0); } while (0)
    const int qt = blockIdx.x % p.QT;
    const int s = blockIdx.x / p.QT;
    const int64_t t_begin = p.tiles_total * s / p.S;
    const int64_t t_end = p.tiles_total * (s + 1) / p.S;

    if (tid == 0) {
        // pair: the leader's full / q / tempty barriers collect both CTAs (one producer arrival and the
        // bytes of each; every epilogue thread of the pair); empty / tfull exist per CTA and are signalled
        // together by the leader's multicast commits
        for (int i = 0; i < stages; i++) {
            mbar_init(&full_bar[i], (uint32_t)p.csize);
            mbar_init(&empty_bar[i], 1);
        }
        mbar_init(q_bar, (uint32_t)p.csize);
        for (int a = 0; a < 2; a++) {
            mbar_init(&tfull_bar[a], 1);
            mbar_init(&tempty_bar[a], TC_EPI_THREADS * (uint32_t)p.csize);
            mbar_init(&nfull_bar[a], 1);
            mbar_init(&tlocal_bar[a], TC_EPI_THREADS);
        }
        fence_mbar_init();
    }
    if (tid < 128) thr_sh[tid] = FLT_MAX;
    for (int i = tid; i < 2 * TC_BN; i += TC_THREADS) bias_s[i] = 0.f;  // inner product: no bias
    if (tid == 0) *done_flag = 0;
    if (warp == 2) {
        if constexpr (PAIR) tmem_alloc2<512>(tmem_slot);
        else tmem_alloc<512>(tmem_slot);
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmX);
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
    uint32_t crank = 0u;
    if constexpr (PAIR) crank = cluster_ctarank();
    const bool leader = crank == 0;
    // Everything above (barriers, TMEM, descriptor prefetch) touched no global memory and may have run while
    // prep_queries_kernel was still finishing (programmatic dependent launch); from here on its outputs
    // (bf16 queries, slots, margins) are read.  K5 may be scheduled onto SMs as CTAs of this grid retire.
    pdl_wait();
    pdl_launch_dependents();
    if (p.tstamp != nullptr && tid == 0) atomicMin(p.tstamp, globaltimer_ns());

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            if (p.a_resident) {
                if constexpr (!PAIR) {
                    mbar_expect_tx(q_bar, (uint32_t)p.KB * TC_A_BLOCK);
                    for (int kb = 0; kb < p.KB; kb++)
                        tma_load_2d(sA + (size_t)kb * TC_A_BLOCK, &tmQ, kb * 64, qt * TC_BM, q_bar);
                } else {
                    if (leader) mbar_expect_tx(q_bar, 2u * (uint32_t)p.KB * TC_A_BLOCK);
                    for (int kb = 0; kb < p.KB; kb++)
                        tma_load_2d_2sm(sA + (size_t)kb * TC_A_BLOCK, &tmQ, kb * 64, qt * TC_BM, q_bar);
                    if (!leader) mbar_arrive_remote(q_bar, 0);
                }
            }
            uint32_t st = 0, ph = 0;
            for (int64_t t = t_begin; t < t_end; t++) {
                for (int kb = 0; kb < p.KB; kb++) {
                    TC_WAIT(&empty_bar[st], ph ^ 1u, 0);
                    if (kb == 0) TC_MARK(0, t - t_begin);
                    if (kb == p.KB - 1) TC_MARK(1, t - t_begin);
                    unsigned char* stage = sB + (size_t)st * stage_bytes;
                    if constexpr (!PAIR) {
                        mbar_expect_tx(&full_bar[st], (uint32_t)stage_bytes);
                        tma_load_2d(stage, &tmX, kb * 64, (int)(t * TC_BN), &full_bar[st]);
                        if (!p.a_resident) tma_load_2d(stage + b_bytes, &tmQ, kb * 64, qt * TC_BM, &full_bar[st]);
                    } else {
                        // each CTA stages its half of the corpus tile (and its own query K-block) in its own
                        // shared memory; all bytes are accounted on the leader's barrier
                        if (leader) mbar_expect_tx(&full_bar[st], 2u * (uint32_t)stage_bytes);
                        tma_load_2d_2sm(stage, &tmX, kb * 64, (int)(t * TC_BN) + (int)crank * (TC_BN / 2), &full_bar[st]);
                        if (!p.a_resident) tma_load_2d_2sm(stage + b_bytes, &tmQ, kb * 64, qt * TC_BM, &full_bar[st]);
                        if (!leader) mbar_arrive_remote(&full_bar[st], 0);
                    }
                    if (++st == (uint32_t)stages) {
                        st = 0;
                        ph ^= 1u;
                    }
                }
            }
            if (TC_DBG) p.dbg[blockIdx.x * 8 + 0] = dbg_acc[0];
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (converged warp, one elected lane issues) =====================
        // This warp paces the tensor pipe: 4 MMAs of 128 tensor-cycles per K-block, so its loop has to issue
        // in well under 512 cycles.  All 32 lanes run the loop together (waits included) and only the
        // tcgen05 instructions are predicated on elect.sync: with a single active lane ptxas wraps every
        // tcgen05.mma in a vote/elect/broadcast loop (ncu: ~125 SASS instructions per K-block, the warp
        // issuing for >50 % of its samples and the epilogue idle a third of the time).
        if (leader) {
            const uint32_t idesc = make_idesc_bf16(pair ? 2 * TC_BM : TC_BM, TC_BN);
            const uint64_t desc_hi = make_smem_desc_sw128(0) & 0xffffffff00000000ull;
            const uint32_t desc_lo0 = (uint32_t)(make_smem_desc_sw128(0) & 0xffffffffull);
            const uint32_t a_lo0 = desc_lo0 + ((smem_u32(sA) & 0x3ffffu) >> 4);
            const uint32_t b_lo0 = desc_lo0 + ((smem_u32(sB) & 0x3ffffu) >> 4);
            const uint32_t stage_lo = (uint32_t)stage_bytes >> 4;
            const uint32_t a_in_stage_lo = (uint32_t)b_bytes >> 4;
            if (p.a_resident) TC_WAIT(q_bar, 0, 3);
            uint32_t st = 0, ph = 0, tcount = 0;
            for (int64_t t = t_begin; t < t_end; t++, tcount++) {
                const uint32_t a = tcount & 1u, aph = (tcount >> 1) & 1u;
                TC_WAIT(&tempty_bar[a], aph ^ 1u, 2);  // every epilogue thread (of the pair) has drained accumulator a
                if (lane == 0) TC_MARK(2, tcount);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + a * TC_BN;
                uint32_t a_lo = a_lo0;  // resident query tile: K-block kb at + kb * 16 KiB
                for (int kb = 0; kb < p.KB; kb++) {
                    TC_WAIT(&full_bar[st], ph, 1);
                    if (lane == 0) {
                        if (kb == 0) TC_MARK(3, tcount);
                        if (kb == p.KB - 1) TC_MARK(4, tcount);
                    }
                    tc_fence_after();
                    const uint32_t b_lo = b_lo0 + st * stage_lo;
                    const uint32_t a_blk = p.a_resident ? a_lo : b_lo + a_in_stage_lo;
                    if (elect_one()) {
#pragma unroll
                        for (int kk = 0; kk < 4; kk++) {
                            const uint64_t da = desc_hi | (uint64_t)(a_blk + 2 * kk);  // + 32 bytes per K = 16 step
                            const uint64_t db = desc_hi | (uint64_t)(b_lo + 2 * kk);
                            const uint32_t acc = (kb | kk) != 0 ? 1u : 0u;
                            if constexpr (PAIR) umma_bf16_2sm(d_tmem, da, db, idesc, acc);
                            else umma_bf16(d_tmem, da, db, idesc, acc);
                        }
                        if constexpr (PAIR) umma_commit_2sm(&empty_bar[st]);  // frees the stage in both CTAs
                        else umma_commit(&empty_bar[st]);           // smem stage reusable once these MMAs retire
                    }
                    __syncwarp();
                    a_lo += TC_A_BLOCK >> 4;
                    if (++st == (uint32_t)stages) {
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
            if (TC_DBG && lane == 0) {
                p.dbg[blockIdx.x * 8 + 1] = dbg_acc[1];
                p.dbg[blockIdx.x * 8 + 2] = dbg_acc[2];
            }
        }
    } else if (warp == 2) {
        // ===================== norms fetcher =====================
        // bias slot a is free once THIS CTA's epilogue has released accumulator a: fetch the tile's 256
        // squared norms then; they land long before the epilogue reaches the tile
        if (lane == 0 && p.use_bias) {
            uint32_t tcount = 0;
            for (int64_t t = t_begin; t < t_end; t++, tcount++) {
                const uint32_t a = tcount & 1u, aph = (tcount >> 1) & 1u;
                mbar_wait_parked(&tlocal_bar[a], aph ^ 1u);
                mbar_expect_tx(&nfull_bar[a], TC_BN * 4);
                bulk_g2s(bias_s + a * TC_BN, p.nrm + t * TC_BN, TC_BN * 4, &nfull_bar[a]);
            }
        }
    } else if (warp == 3) {
        // ===================== threshold helper =====================
        // keeps thr_sh[row] = max over the query's KP published slots (an upper bound on its global
        // KP-th best key) fresh, so the epilogue never waits on global memory for it
        const unsigned int unset = f2ord(FLT_MAX);
        while (*done_flag == 0) {
#pragma unroll 1
            for (int rr = 0; rr < 4; rr++) {
                const int row = rr * 32 + lane;
                const int qrow = qt * TC_BM + row;
                if (qrow < p.nq) {
                    const uint4* sl = reinterpret_cast<const uint4*>(p.slots + (size_t)qrow * p.NS);
                    unsigned int m = 0;
#pragma unroll 4
                    for (int i = 0; i < p.NS / 4; i++) {
                        uint4 v;
                        asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];"
                                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                                     : "l"(sl + i));
                        m = max(max(m, v.x), max(max(v.y, v.z), v.w));
                    }
                    if (m < unset) thr_sh[row] = ord2f(m);
                }
            }
            __nanosleep(1000);
        }
    } else if (warp >= 4) {
        // ============ epilogue: 8 warps; thread == (query row, column half of the tile) ============
        const int et = tid - 128;          // 0..255
        const int half = et >> 7;          // which 128 columns of each 256-column tile
        const int row = et & 127;          // query row in the tile == TMEM lane
        const int qrow = qt * TC_BM + row;
        const bool active = qrow < p.nq;
        float skey[KP];
        uint32_t sidx[KP];
#pragma unroll
        for (int i = 0; i < KP; i++) {
            skey[i] = FLT_MAX;
            sidx[i] = 0xffffffffu;
        }
        // effective threshold = min(own KP-th best, best KP-th any thread has published for this
        // query + the query's error margin): rows above it can be neither in this thread's top-KP nor
        // within the margin band of the global top-KP, which is all K5's certificate needs
        const float qmargin = active ? p.qmargin[qrow] : 0.f;
        unsigned int* my_slots = p.slots + (size_t)(active ? qrow : 0) * p.NS;
        const int grp = (s * 2 + half) & (p.G - 1);
        float thr = active ? FLT_MAX : -FLT_MAX;
        int cnt = 0;
        uint64_t* my_pend = (p.pend_g ? p.pend_g + (size_t)blockIdx.x * TC_BUFCAP * TC_EPI_THREADS : pend_s) + et;  // [slot][thread]
        const float scale = p.is_l2 ? -2.0f : -1.0f;

        // sorted insert without a dependency chain: every slot decides from the OLD neighbours
        auto insert = [&](float x, uint32_t xi) {
            bool cj = x < skey[KP - 1];
#pragma unroll
            for (int j = KP - 1; j >= 1; j--) {
                const bool cjm1 = x < skey[j - 1];
                skey[j] = cjm1 ? skey[j - 1] : (cj ? x : skey[j]);
                sidx[j] = cjm1 ? sidx[j - 1] : (cj ? xi : sidx[j]);
                cj = cjm1;
            }
            skey[0] = cj ? x : skey[0];
            sidx[0] = cj ? xi : sidx[0];
        };
        auto refresh_thr = [&]() {
            if (active) {
                const float g = *reinterpret_cast<volatile float*>(thr_sh + row);
                // both terms only fall over time; the min with the old value keeps the tile-0 bound until they pass it
                thr = fminf(thr, fminf(skey[KP - 1], g < FLT_MAX ? g + qmargin : FLT_MAX));
            }
        };
        auto flush = [&]() {
            const int maxc = __reduce_max_sync(0xffffffffu, cnt);
#pragma unroll 1
            for (int i = 0; i < maxc; i++) {
                const bool has = i < cnt;
                const uint64_t e = my_pend[(size_t)i * TC_EPI_THREADS];
                const float x = __uint_as_float((uint32_t)(e >> 32));
                if (has && x < skey[KP - 1]) insert(x, (uint32_t)e);
            }
            cnt = 0;
            if (active) {
#pragma unroll
                for (int j = 0; j < KP; j++)
                    if (j < p.J && skey[j] < FLT_MAX) atomicMin(my_slots + grp * p.J + j, f2ord(skey[j]));
            }
            refresh_thr();
        };

        uint32_t tcount = 0;
        for (int64_t t = t_begin; t < t_end; t++, tcount++) {
            const uint32_t a = tcount & 1u, aph = (tcount >> 1) & 1u;
            refresh_thr();
            {
                const long long _t0 = TC_DBG ? clock64() : 0;
                mbar_wait(&tfull_bar[a], aph);  // accumulator complete
                if (p.use_bias) mbar_wait(&nfull_bar[a], aph);  // and the tile's norms / selector bias have landed
                if (TC_DBG) dbg_acc[0] += (unsigned long long)(clock64() - _t0);
                if (et == 0) TC_MARK(5, tcount);
            }
            if ((t + 1) * TC_BN > p.n) {  // ragged last tile of the corpus: columns past the end never win
                if (t * TC_BN + et >= p.n) bias_s[a * TC_BN + et] = INFINITY;
                epi_sync();
            }
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(((warp - 4) & 3) * 32) << 16) + a * TC_BN + half * 128;
            const uint32_t col0 = (uint32_t)(t * TC_BN) + half * 128;
            const float* bias = bias_s + a * TC_BN + half * 128;
            if (tcount == 0 && p.prepass) {
                // Tile 0, first reading: only the thread's smallest key, published to its group's slot.  With
                // J == 1 a slot is the smallest key any stream of the group has seen, so once the G >= k slots of
                // a query are set, max(slots) bounds its k-th best key -- from 256 rows of EVERY stream of the
                // grid rather than this thread's own 128.  Without it every thread starts at +inf: its first 16
                // columns all survive, the list only learns over the tile, and with 32 different queries per
                // warp nearly every group of tile 0 walks the slow path (49 appends and 4 flushes per thread).
                float mn = FLT_MAX;
                uint32_t r0[16];
#pragma unroll 1
                for (int sc = 0; sc < 8; sc++) {
                    tmem_ld_32x16(taddr + sc * 16, r0);
                    tmem_ld_wait();
                    const float4* b4 = reinterpret_cast<const float4*>(bias + sc * 16);
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++) {
                        const float4 b = b4[j4];
                        mn = fminf(mn, fminf(fminf(fmaf(scale, __uint_as_float(r0[4 * j4 + 0]), b.x),
                                                   fmaf(scale, __uint_as_float(r0[4 * j4 + 1]), b.y)),
                                             fminf(fmaf(scale, __uint_as_float(r0[4 * j4 + 2]), b.z),
                                                   fmaf(scale, __uint_as_float(r0[4 * j4 + 3]), b.w))));
                    }
                }
                if (active && mn < FLT_MAX) atomicMin(my_slots + grp, f2ord(mn));
                // the other streams reach this point within a microsecond or two of each other (every CTA's
                // first tile comes out of the same HBM burst): poll the query's slots for a bounded time
                const unsigned int unset = f2ord(FLT_MAX);
                const long long t0 = clock64();
                bool got = !active;
                for (;;) {
                    if (!got) {
                        const uint4* sl = reinterpret_cast<const uint4*>(my_slots);
                        unsigned int m = 0;
                        for (int i = 0; i < p.NS / 4; i++) {
                            uint4 v;
                            asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];"
                                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                                         : "l"(sl + i));
                            m = max(max(m, v.x), max(max(v.y, v.z), v.w));
                        }
                        if (m < unset) {
                            thr = fminf(thr, ord2f(m) + qmargin);
                            got = true;
                        }
                    }
                    const bool more = !got && (clock64() - t0 < 6000);  // ~3 us, then carry on without the bound
                    if (!__any_sync(0xffffffffu, more)) break;
                }
            }
            // 16 columns at a time, kept as a rolled loop (two register sets, next load in flight while
            // the current set is filtered): the fully unrolled form of this loop was 58 KB of SASS and
            // stalled on instruction fetch
            const uint64_t scale2 = f2_pack(scale, scale);
            // NOBIAS (inner product, no selector, not the ragged last tile): the key is just -dot, so the 16 FMAs and
            // the four bias loads of a group are not needed to find out that it holds no survivor
            auto process16 = [&](const uint32_t (&r)[16], int c0, auto nobias) {
                constexpr bool NOBIAS = decltype(nobias)::value;
                float kk[16], m4[4];  // keys, and per-quad minima: the slow path below only touches quads that hold a survivor
                if constexpr (NOBIAS) {
                    // min of the keys of a quad = -(max of its dots); the keys themselves (0 - dot, the value the FMA
                    // with a zero bias gives, signed zeros included) are only formed where they are stored
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++)
                        m4[j4] = 0.f - fmaxf(fmaxf(__uint_as_float(r[4 * j4 + 0]), __uint_as_float(r[4 * j4 + 1])),
                                             fmaxf(__uint_as_float(r[4 * j4 + 2]), __uint_as_float(r[4 * j4 + 3])));
#pragma unroll
                    for (int i = 0; i < 16; i++) kk[i] = 0.f - __uint_as_float(r[i]);
                } else {
                    const float4* b4 = reinterpret_cast<const float4*>(bias + c0);
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++) {  // key = scale * dot + bias, two columns per FFMA2 (same rounding as fmaf)
                        const float4 b = b4[j4];
                        f2_unpack(f2_fma(scale2, f2_pack_bits(r[4 * j4 + 0], r[4 * j4 + 1]), f2_pack(b.x, b.y)), kk[4 * j4 + 0], kk[4 * j4 + 1]);
                        f2_unpack(f2_fma(scale2, f2_pack_bits(r[4 * j4 + 2], r[4 * j4 + 3]), f2_pack(b.z, b.w)), kk[4 * j4 + 2], kk[4 * j4 + 3]);
                    }
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++)
                        m4[j4] = fminf(fminf(kk[4 * j4 + 0], kk[4 * j4 + 1]), fminf(kk[4 * j4 + 2], kk[4 * j4 + 3]));
                }
                const float kmin = fminf(fminf(m4[0], m4[1]), fminf(m4[2], m4[3]));
                // after the first tiles most groups of 16 have no survivor in any lane.  The lanes of a warp
                // are 32 different queries with 32 different thresholds, so once one lane has a survivor the
                // whole warp walks the slow path: it is kept to the quads that are hit (one, typically)
                if (__any_sync(0xffffffffu, kmin < thr)) {
                    unsigned int qm = 0;
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++) qm |= (m4[j4] < thr) ? (1u << j4) : 0u;
                    // flush only when this group's survivors might not fit behind the pending ones, so one
                    // flush serves many groups (thr can only fall inside: qm stays a superset)
                    if (__any_sync(0xffffffffu, cnt + 4 * __popc(qm) > TC_BUFCAP)) {
                        const long long _t0 = TC_DBG ? clock64() : 0;
                        if (DBG) dbg_acc[2] += (unsigned long long)cnt;
                        flush();
                        if (TC_DBG) {
                            dbg_acc[1] += (unsigned long long)(clock64() - _t0);
                            dbg_acc[3] += 1;
                        }
                    }
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++) {
                        if (__any_sync(0xffffffffu, (qm >> j4) & 1u)) {
#pragma unroll
                            for (int j = 0; j < 4; j++) {
                                if (kk[4 * j4 + j] < thr) {
                                    my_pend[(size_t)cnt * TC_EPI_THREADS] =
                                        ((uint64_t)__float_as_uint(kk[4 * j4 + j]) << 32) | (uint64_t)(col0 + c0 + 4 * j4 + j);
                                    cnt++;
                                }
                            }
                        }
                    }
                }
            };
            uint32_t ra[16], rb[16];
            tmem_ld_32x16(taddr, ra);
            if (!p.is_l2 && !p.use_bias && (t + 1) * TC_BN <= p.n) {
#pragma unroll 1
                for (int sc = 0; sc < 8; sc += 2) {
                    tmem_ld_wait();
                    tmem_ld_32x16(taddr + (sc + 1) * 16, rb);
                    process16(ra, sc * 16, std::true_type());
                    tmem_ld_wait();
                    if (sc + 2 < 8) tmem_ld_32x16(taddr + (sc + 2) * 16, ra);
                    process16(rb, (sc + 1) * 16, std::true_type());
                }
            } else {
#pragma unroll 1
                for (int sc = 0; sc < 8; sc += 2) {
                    tmem_ld_wait();
                    tmem_ld_32x16(taddr + (sc + 1) * 16, rb);
                    process16(ra, sc * 16, std::false_type());
                    tmem_ld_wait();
                    if (sc + 2 < 8) tmem_ld_32x16(taddr + (sc + 2) * 16, ra);
                    process16(rb, (sc + 1) * 16, std::false_type());
                }
            }
            tc_fence_before();
            mbar_arrive(&tlocal_bar[a]);
            if (leader) mbar_arrive(&tempty_bar[a]);
            else if constexpr (PAIR) mbar_arrive_remote(&tempty_bar[a], 0);  // the MMA issuer lives in the leader CTA
            if (et == 0) TC_MARK(6, tcount);
        }
        if (DBG) dbg_acc[2] += (unsigned long long)cnt;
        flush();
        if (et == 0) *done_flag = 1;
        if (TC_DBG && et == 0) {
            p.dbg[blockIdx.x * 8 + 3] = dbg_acc[0];                         // epilogue: waiting for an accumulator
            p.dbg[blockIdx.x * 8 + 4] = dbg_acc[1];                         // epilogue: cycles inside flush()
            p.dbg[blockIdx.x * 8 + 5] = dbg_acc[2];                         // entries appended by thread 0
            p.dbg[blockIdx.x * 8 + 6] = dbg_acc[3];                         // flushes of warp 4
            p.dbg[blockIdx.x * 8 + 7] = (unsigned long long)(clock64() - dbg_t0);  // CTA lifetime
        }
        // ---- every thread-stream (query, stream, column half) publishes its own list + drop threshold ----
        if (active) {
            const size_t ts = (size_t)qrow * (2 * p.S) + (size_t)s * 2 + half;
            uint64_t* dst = p.out_cand + ts * KP;
#pragma unroll
            for (int i = 0; i < KP; i++) dst[i] = (sidx[i] == 0xffffffffu) ? COMP_EMPTY : make_comp(skey[i], sidx[i]);
            // every row this thread dropped had key >= the threshold in force at that moment, and
            // thresholds only fall
            p.out_thr[ts] = fminf(thr, skey[KP - 1]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // nobody leaves while the peer may still signal its barriers
    if (warp == 2) {
        tc_fence_after();
        if constexpr (PAIR) tmem_dealloc2<512>(tmem_base);
        else tmem_dealloc<512>(tmem_base);
    }
    if (p.tstamp != nullptr && tid == 0) atomicMax(p.tstamp + 1, globaltimer_ns());
}

// queries fp32 [nq, ldq] -> bf16 [nq_pad, ld16] (zero padded rows and columns) + ||q||^2
__global__ void __launch_bounds__(256) prep_queries_kernel(const float* __restrict__ q, int ldq, int d, int64_t nq,
                                                           int64_t nq_pad, __nv_bfloat16* __restrict__ q16, int ld16,
                                                           float* __restrict__ qn, const unsigned int* __restrict__ maxn,
                                                           float c_acc, float c_abs, int is_l2, float* __restrict__ qmargin,
                                                           unsigned int* __restrict__ slots, int NS,
                                                           int* __restrict__ flags, unsigned long long* tstamp,
                                                           float* __restrict__ q_copy, int ld_copy) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    pdl_launch_dependents();  // the scan's prologue may start now; it waits for this grid before reading anything
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (tstamp) atomicMin(tstamp + 2, globaltimer_ns());
        flags[0] = 0;  // uncertified-query count of this pass (K5 appends)
        flags[-1] = 0; // K5's CTA ticket (the last CTA to finish publishes the count to the host)
    }
    if (row >= nq_pad) return;
    float ss = 0.f, sh = 0.f, se = 0.f;  // ||q||^2, ||bf16 q||^2, ||q - bf16 q||^2
    auto take = [&](float v, int c) {
        const __nv_bfloat16 b = __float2bfloat16_rn(v);
        const float vb = __bfloat162float(b);
        ss += v * v;
        sh += vb * vb;
        se += (v - vb) * (v - vb);
        q16[row * ld16 + c] = b;
        // the queries may sit in pinned HOST memory (read here once, over PCIe): K5 and the fall-back scan get a
        // zero-padded fp32 copy in device memory
        if (q_copy != nullptr && row < nq && c < ld_copy) q_copy[row * ld_copy + c] = v;
    };
    const float* src = q + row * ldq;
    if ((d & 3) == 0 && (ldq & 3) == 0 && (reinterpret_cast<uintptr_t>(q) & 15u) == 0) {
        // 16-byte loads, four per lane issued before the first is used: a row is one or two round trips to wherever
        // the queries live (over PCIe when that is pinned host memory), not ld16 / 32 of them
        for (int c0 = 4 * lane; c0 < ld16; c0 += 512) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int c = c0 + 128 * u;
                v[u] = (row < nq && c < d) ? *reinterpret_cast<const float4*>(src + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int c = c0 + 128 * u;
                if (c < ld16) {
                    take(v[u].x, c);
                    take(v[u].y, c + 1);
                    take(v[u].z, c + 2);
                    take(v[u].w, c + 3);
                }
            }
        }
    } else {
        for (int c = lane; c < ld16; c += 32) take((row < nq && c < d) ? src[c] : 0.f, c);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ss += __shfl_xor_sync(0xffffffffu, ss, o);
        sh += __shfl_xor_sync(0xffffffffu, sh, o);
        se += __shfl_xor_sync(0xffffffffu, se, o);
    }
    if (lane == 0) {
        qn[row] = ss;
        // eps bounds |<q,x> - acc| for every row x, acc = the tensor pipe's value of <bf16 q, xs> over the scan copy xs:
        //   <q,x> - <bf16 q, xs> = <q - bf16 q, x> + <bf16 q, x - xs>     (Cauchy-Schwarz on both, with the MEASURED
        //   rounding losses ||q - bf16 q|| of this query and max ||x - xs|| of the corpus instead of the worst case
        //   2^-9 per element: about 1.7x tighter, and nothing at all for a bf16-stored corpus, where xs = x)
        //   + c_acc ||bf16 q|| ||xs||  for the pipe's fp32 accumulation (see tc_pass).
        // 2*eps in key space: L2 keys carry -2<q,x>, so 4*eps; inner-product keys 2*eps.  1.001: the norms above
        // are themselves fp32 sums.
        const float mx = __uint_as_float(maxn[0]), me = __uint_as_float(maxn[1]);
        const float eps = 1.001f * (sqrtf(se * mx) + sqrtf(sh * me) + c_acc * sqrtf(sh * mx) * 1.004f);
        qmargin[row] = (is_l2 ? 4.f : 2.f) * eps + c_abs * (ss + mx);
    }
    for (int i = lane; i < NS; i += 32) slots[row * NS + i] = f2ord(FLT_MAX);
}

struct RerankParams {
    const uint64_t* out_cand;  // [nq_pad][S][KP]
    const float* out_thr;      // [nq_pad][S]
    const float* qmargin;      // [nq_pad] 2*eps in key space
    const float* q;            // [nq][ldq] fp32
    int ldq;
    int S;
    int KP;
    int k;
    int is_l2;
    int storage;
    const float* x32;
    int ld32;
    const __nv_bfloat16* x16;
    int ld16;
    int d;
    int64_t id_offset;
    float* D;                  // [nq][k]
    int64_t* I;
    int* nflag;       // [-1] = CTA ticket, [0] = count, [1..] = list
    int* flag_list;
    int out_is_host;  // D / I point into pinned host memory: results must be visible system-wide before the ticket
    int* h_nflag;     // pinned host words (device-mapped): [0] the count, [1] `seq`, written in that order by the last CTA
    int seq;
    unsigned long long* dbg;  // optional [16] phase clocks of CTA 0 (BFK_TC_DEBUG), else nullptr
    unsigned long long* tstamp;  // optional: [3] = latest CTA end (globaltimer ns), see TcParams
};

// Sorts a[0..n) ascending (n a power of two); called by every thread of the CTA, results visible on return.
// Up to 64 * (warps in the CTA) elements, warp w keeps elements [64w, 64w + 64) in registers (two per lane):
// compare-exchange partners i ^ j are another lane's same slot (shuffle) for j < 32, this lane's other slot for
// j = 32, and only j >= 64 goes through shared memory with a CTA barrier -- 3 barrier steps at n = 256 where the
// all-shared-memory network needs 36 (it was a third of K5's time at nq = 64).
__device__ void rr_bitonic(uint64_t* a, int n, int tid, int nthr) {
    const int lane = tid & 31, warp = tid >> 5;
    if (n <= 64 * (nthr >> 5)) {
        __syncthreads();
        const int i0 = warp * 64 + lane, i1 = i0 + 32;
        const bool mine = warp * 64 < n;
        uint64_t v0 = (mine && i0 < n) ? a[i0] : COMP_EMPTY;
        uint64_t v1 = (mine && i1 < n) ? a[i1] : COMP_EMPTY;
        for (int k2 = 2; k2 <= n; k2 <<= 1) {
            const bool up0 = (i0 & k2) == 0, up1 = (i1 & k2) == 0;
            for (int j = k2 >> 1; j > 0; j >>= 1) {
                if (j >= 64) {
                    if (mine) {
                        a[i0] = v0;
                        a[i1] = v1;
                    }
                    __syncthreads();
                    if (mine) {
                        const uint64_t o0 = a[i0 ^ j], o1 = a[i1 ^ j];
                        const bool lower = (i0 & j) == 0;  // bit j >= 64 is the same for both slots
                        v0 = (lower == up0) ? (v0 < o0 ? v0 : o0) : (v0 < o0 ? o0 : v0);
                        v1 = (lower == up1) ? (v1 < o1 ? v1 : o1) : (v1 < o1 ? o1 : v1);
                    }
                    __syncthreads();
                } else if (j == 32) {
                    const uint64_t lo = v0 < v1 ? v0 : v1, hi = v0 < v1 ? v1 : v0;
                    v0 = up0 ? lo : hi;  // k2 >= 64 here: both slots sort in the same direction
                    v1 = up0 ? hi : lo;
                } else {
                    const bool lower = (lane & j) == 0;
                    const uint64_t o0 = __shfl_xor_sync(0xffffffffu, v0, j);
                    const uint64_t o1 = __shfl_xor_sync(0xffffffffu, v1, j);
                    v0 = (lower == up0) ? (v0 < o0 ? v0 : o0) : (v0 < o0 ? o0 : v0);
                    v1 = (lower == up1) ? (v1 < o1 ? v1 : o1) : (v1 < o1 ? o1 : v1);
                }
            }
        }
        if (mine && i0 < n) a[i0] = v0;
        if (mine && i1 < n) a[i1] = v1;
        __syncthreads();
        return;
    }
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

// Sorts a[0..n) ascending, n a power of two.  Up to 256 elements every element's rank is COUNTED: n*n
// independent compares spread over all threads (NT / n threads per element, each looking at every (NT/n)-th
// element), one shuffle reduction, one scatter.  The bitonic network on so few elements is a single warp
// walking a chain of 21-36 dependent exchange steps -- 3 us each time, twice per query, a quarter of K5.
// Equal composites (only COMP_EMPTY padding repeats) are ordered by position so that ranks stay distinct.
template <int NT>
__device__ void rr_sort(uint64_t* a, int n, int tid) {
    if (n > 256 || n > NT) {
        rr_bitonic(a, n, tid, NT);
        return;
    }
    __syncthreads();
    const int P = NT / n;  // power of two, <= 16: the P threads of an element are neighbours in one warp
    const int e = tid / P, part = tid - e * P;
    const uint64_t v = a[e];
    int cnt = 0;
    for (int j = part; j < n; j += P) {
        const uint64_t o = a[j];
        cnt += (o < v || (o == v && j < e)) ? 1 : 0;
    }
    for (int o = P >> 1; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    __syncthreads();
    if (part == 0) a[cnt] = v;
    __syncthreads();
}

// RR_THREADS = 512 for long candidate lists (few queries, many streams), 128 when a query has at most 1024
// candidates (many queries): four times as many queries are then resident per SM.
template <int RR_THREADS>
__global__ void __launch_bounds__(RR_THREADS) rerank_kernel(const RerankParams p) {
    // the 128-thread form only ever sees <= 1024 candidates per query: an 8 KB window instead of 32 KB (five CTAs per
    // SM instead of four; measured no faster -- at 2048 queries K5 is bound by the ~60 MB of scattered row reads)
    constexpr int MERGE_N = RR_THREADS == 128 ? 1024 : TC_MERGE_N;
    __shared__ uint64_t buf[MERGE_N];
    __shared__ uint64_t ex[TC_RMAX];
    __shared__ float q_s[BFK_MAX_D];
    __shared__ float red[RR_THREADS / 32];
    __shared__ int info[4];
    __shared__ int cnt_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t q = blockIdx.x;
    const int64_t M = (int64_t)p.S * p.KP;
    const uint64_t* src = p.out_cand + q * M;
    pdl_wait();  // launched behind the scan with programmatic serialization: its lists are complete from here on
    const long long rr_t0 = p.dbg ? clock64() : 0;
    const unsigned long long rr_g0 = p.dbg ? globaltimer_ns() : 0ull;
#define RR_MARK(i) do { if (p.dbg && blockIdx.x == 0 && tid == 0) p.dbg[i] = (unsigned long long)(clock64() - rr_t0); } while (0)

    // two candidates per 16-byte load; the first four loads of every thread are issued now, together with the
    // thresholds and the query, so that the filter below does not start a second round trip to L2
    // (KP is 16 or 32, so a query's M = S * KP candidates are an even count at a 16-byte aligned address)
    const ulonglong2* src2 = reinterpret_cast<const ulonglong2*>(src);
    const int64_t M2 = M >> 1;
    ulonglong2 pre[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int64_t i = tid + (int64_t)u * RR_THREADS;
        pre[u] = i < M2 ? __ldcg(src2 + i) : make_ulonglong2(COMP_EMPTY, COMP_EMPTY);
    }
    for (int c = tid; c < max(p.ld32, p.ld16); c += RR_THREADS) q_s[c] = (c < p.ldq) ? p.q[q * p.ldq + c] : 0.f;  // zero padded: the row loads below run to the padded width
    // T = min over streams of the drop threshold: every row some stream dropped has approximate key >= T
    float tmin = FLT_MAX;
    for (int s = tid; s < p.S; s += RR_THREADS) tmin = fminf(tmin, p.out_thr[q * p.S + s]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tmin = fminf(tmin, __shfl_xor_sync(0xffffffffu, tmin, o));
    if (lane == 0) red[warp] = tmin;
    if (tid == 0) cnt_s = 0;
    __syncthreads();
    tmin = red[0];
#pragma unroll
    for (int w = 1; w < RR_THREADS / 32; w++) tmin = fminf(tmin, red[w]);
    RR_MARK(0);

    // ---- keep only candidates with approximate key <= T (a certified answer never needs the rest:
    //      the cut A_k + 2*eps must lie below T), then sort the few survivors ----
    {
        const uint32_t tmin_ord = f2ord(tmin);
        auto keep = [&](uint64_t c) {
            if (c != COMP_EMPTY && (uint32_t)(c >> 32) <= tmin_ord) {
                const int slot = atomicAdd(&cnt_s, 1);
                if (slot < MERGE_N) buf[slot] = c;
            }
        };
#pragma unroll
        for (int u = 0; u < 4; u++) {
            keep(pre[u].x);
            keep(pre[u].y);
        }
        for (int64_t i0 = tid + 4 * (int64_t)RR_THREADS; i0 < M2; i0 += 4 * RR_THREADS) {
            ulonglong2 v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int64_t i = i0 + (int64_t)u * RR_THREADS;
                v[u] = i < M2 ? __ldcg(src2 + i) : make_ulonglong2(COMP_EMPTY, COMP_EMPTY);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                keep(v[u].x);
                keep(v[u].y);
            }
        }
    }
    __syncthreads();
    RR_MARK(1);
    const int total = cnt_s;
    const int cnt = min(total, MERGE_N);
    int n = 64;
    while (n < cnt) n <<= 1;
    for (int i = cnt + tid; i < n; i += RR_THREADS) buf[i] = COMP_EMPTY;
    __syncthreads();
    rr_sort<RR_THREADS>(buf, n, tid);
    RR_MARK(2);
    if (p.dbg && blockIdx.x == 0 && tid == 0) p.dbg[8] = (unsigned long long)total;

    // ---- the cut: A_k + 2*eps in approximate-key space (buf is sorted: the rows inside the cut are a prefix,
    //      counted by all threads at once) ----
    {
        const int nvalid = min(cnt, TC_RMAX);
        const int kk = min(p.k, nvalid);
        float cut = 0.f;
        int inside = 0;
        if (nvalid > 0) {
            cut = ord2f((uint32_t)(buf[kk - 1] >> 32)) + p.qmargin[q];
            for (int i = tid; i < nvalid; i += RR_THREADS) inside += ord2f((uint32_t)(buf[i] >> 32)) <= cut ? 1 : 0;
        }
        if (tid == 0) info[2] = 0;
        __syncthreads();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) inside += __shfl_xor_sync(0xffffffffu, inside, o);
        if (lane == 0 && inside) atomicAdd(&info[2], inside);
        __syncthreads();
        int R = info[2];
        if (tid == 0) {
            int certified = 1;
            if (nvalid > 0) {
                if (R < kk) R = kk;
                if (!(tmin > cut)) certified = 0;          // a dropped row could still reach the cut
                if (total > MERGE_N) certified = 0;        // survivors did not fit the sort window
                if (R == TC_RMAX && cnt > TC_RMAX && ord2f((uint32_t)(buf[TC_RMAX] >> 32)) <= cut)
                    certified = 0;                         // more rows inside the cut than we re-score
            } else {
                R = 0;
            }
            info[0] = R;
            info[1] = certified;
        }
    }
    for (int i = tid; i < TC_RMAX; i += RR_THREADS) ex[i] = COMP_EMPTY;
    __syncthreads();
    const int R = info[0];
    RR_MARK(3);
    if (p.dbg && blockIdx.x == 0 && tid == 0) p.dbg[9] = (unsigned long long)R;
    int n_ex = 32;
    while (n_ex < R || n_ex < p.k) n_ex <<= 1;  // <= TC_RMAX since R, k <= TC_RMAX

    // ---- exact re-score of the R survivors (faiss's direct-difference arithmetic); a warp works on two rows at
    //      a time so that both rows' (random, DRAM-latency) loads are in flight together ----
    constexpr int NW = RR_THREADS / 32;
    for (int i = warp; i < R; i += 2 * NW) {
        const int i2 = i + NW;
        const bool two = i2 < R;
        const uint32_t id1 = (uint32_t)(buf[i] & 0xffffffffu);
        const uint32_t id2 = two ? (uint32_t)(buf[i2] & 0xffffffffu) : id1;
        float acc1 = 0.f, acc2 = 0.f;
        // All loads of both rows are issued before the first is used (up to four 16-byte loads per row and lane:
        // d <= 512 in fp32, d <= 1024 in bf16 in one go): written as load-use-load-use the three column steps of a
        // 384-wide row were three DRAM latencies in a row, 2.5 us of an 8 us kernel.
        if (p.storage == BFK_STORE_F32) {
            const float4* row1 = reinterpret_cast<const float4*>(p.x32 + (size_t)id1 * p.ld32);
            const float4* row2 = reinterpret_cast<const float4*>(p.x32 + (size_t)id2 * p.ld32);
            const int nv = p.ld32 >> 2;
            for (int c0 = lane; c0 < nv; c0 += 128) {
                float4 a[4], b[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int c = c0 + 32 * u;
                    if (c < nv) {
                        a[u] = __ldg(row1 + c);
                        b[u] = __ldg(row2 + c);
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int c = c0 + 32 * u;
                    if (c < nv) {
                        const float4 x1 = a[u], x2 = b[u];
                        const float4 y = *reinterpret_cast<const float4*>(q_s + 4 * c);
                        if (p.is_l2) {
                            float t0 = x1.x - y.x, t1 = x1.y - y.y, t2 = x1.z - y.z, t3 = x1.w - y.w;
                            acc1 = fmaf(t0, t0, acc1); acc1 = fmaf(t1, t1, acc1); acc1 = fmaf(t2, t2, acc1); acc1 = fmaf(t3, t3, acc1);
                            t0 = x2.x - y.x; t1 = x2.y - y.y; t2 = x2.z - y.z; t3 = x2.w - y.w;
                            acc2 = fmaf(t0, t0, acc2); acc2 = fmaf(t1, t1, acc2); acc2 = fmaf(t2, t2, acc2); acc2 = fmaf(t3, t3, acc2);
                        } else {
                            acc1 = fmaf(x1.x, y.x, acc1); acc1 = fmaf(x1.y, y.y, acc1); acc1 = fmaf(x1.z, y.z, acc1); acc1 = fmaf(x1.w, y.w, acc1);
                            acc2 = fmaf(x2.x, y.x, acc2); acc2 = fmaf(x2.y, y.y, acc2); acc2 = fmaf(x2.z, y.z, acc2); acc2 = fmaf(x2.w, y.w, acc2);
                        }
                    }
                }
            }
        } else {
            // bf16-stored rows: eight elements per 16-byte load (rows are padded with zeros to 64 columns, q_s likewise)
            const uint4* row1 = reinterpret_cast<const uint4*>(p.x16 + (size_t)id1 * p.ld16);
            const uint4* row2 = reinterpret_cast<const uint4*>(p.x16 + (size_t)id2 * p.ld16);
            const int nv = p.ld16 >> 3;
            auto fold = [&](const uint4& v, const float* y, float& acc) {
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float lo = __uint_as_float(w[e] << 16), hi = __uint_as_float(w[e] & 0xffff0000u);
                    if (p.is_l2) {
                        const float t0 = lo - y[2 * e], t1 = hi - y[2 * e + 1];
                        acc = fmaf(t0, t0, acc);
                        acc = fmaf(t1, t1, acc);
                    } else {
                        acc = fmaf(lo, y[2 * e], acc);
                        acc = fmaf(hi, y[2 * e + 1], acc);
                    }
                }
            };
            for (int c0 = lane; c0 < nv; c0 += 128) {
                uint4 a[4], b[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int c = c0 + 32 * u;
                    if (c < nv) {
                        a[u] = __ldg(row1 + c);
                        b[u] = __ldg(row2 + c);
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int c = c0 + 32 * u;
                    if (c < nv) {
                        fold(a[u], q_s + 8 * c, acc1);
                        fold(b[u], q_s + 8 * c, acc2);
                    }
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            acc1 += __shfl_xor_sync(0xffffffffu, acc1, o);
            acc2 += __shfl_xor_sync(0xffffffffu, acc2, o);
        }
        if (lane == 0) {
            const float key1 = p.is_l2 ? acc1 : -acc1;
            ex[i] = (key1 < FLT_MAX) ? make_comp(key1, id1) : COMP_EMPTY;
            if (two) {
                const float key2 = p.is_l2 ? acc2 : -acc2;
                ex[i2] = (key2 < FLT_MAX) ? make_comp(key2, id2) : COMP_EMPTY;
            }
        }
    }
    __syncthreads();
    RR_MARK(4);
    rr_sort<RR_THREADS>(ex, min(n_ex, TC_RMAX), tid);
    RR_MARK(5);
    for (int i = tid; i < p.k; i += RR_THREADS) {
        const uint64_t c = (i < TC_RMAX) ? ex[i] : COMP_EMPTY;
        float dv;
        int64_t iv;
        if (c == COMP_EMPTY) {
            dv = p.is_l2 ? FLT_MAX : -FLT_MAX;
            iv = -1;
        } else {
            const float key = ord2f((uint32_t)(c >> 32));
            dv = p.is_l2 ? key : -key;
            iv = (int64_t)(c & 0xffffffffu) + p.id_offset;
        }
        p.D[q * p.k + i] = dv;
        p.I[q * p.k + i] = iv;
    }
    if (tid == 0 && !info[1]) {
        const int slot = atomicAdd(p.nflag, 1);
        p.flag_list[slot] = (int)q;
    }
    if (p.dbg && tid == 0) {  // per query: CTA lifetime in cycles, rows re-scored, survivors of the filter, start time
        p.dbg[16 + 4 * q + 0] = (unsigned long long)(clock64() - rr_t0);
        p.dbg[16 + 4 * q + 1] = (unsigned long long)R;
        p.dbg[16 + 4 * q + 2] = (unsigned long long)total;
        p.dbg[16 + 4 * q + 3] = rr_g0;
        if (q == 0) p.dbg[10] = (unsigned long long)__float_as_uint(p.qmargin[0]);
    }
    // the last CTA to get here hands the count to the host without a separate D2H copy
    if (p.out_is_host) {  // every thread's (D, I) stores have to reach host memory before this CTA's ticket counts
        __threadfence_system();
        __syncthreads();
    }
    if (tid == 0) {
        if (p.tstamp) atomicMax(p.tstamp + 3, globaltimer_ns());
        __threadfence();
        if (atomicAdd(p.nflag - 1, 1) == (int)gridDim.x - 1) {
            __threadfence();
            *reinterpret_cast<volatile int*>(p.h_nflag) = *reinterpret_cast<volatile int*>(p.nflag);
            __threadfence_system();
            // every CTA's (D, I) rows were written and fenced before its ticket: the host may stop waiting here
            *reinterpret_cast<volatile int*>(p.h_nflag + 1) = p.seq;
            __threadfence_system();
        }
    }
}

// ---- host side -------------------------------------------------------------------------
EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        else
            cudaGetLastError();
    }
    return fn;
}

// bf16 matrix [rows, ld] row-major; box = 64 columns x box_rows, 128-byte swizzle, OOB rows read as 0
int make_tmap(CUtensorMap* m, const void* base, int64_t rows, int ld, int box_rows) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return fail("cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t gdim[2] = {(cuuint64_t)ld, (cuuint64_t)std::max<int64_t>(rows, 1)};
    cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return 0;
}

int make_tmap_head(CUtensorMap* m, const void* base, int64_t rows, int ld, int box_rows) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return fail("cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t gdim[2] = {(cuuint64_t)ld, (cuuint64_t)std::max<int64_t>(rows, 1)};
    cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return 0;
}

// The two tensor maps of a pass change only when the storage is regrown, rows are appended or the batch size
// changes: they are re-encoded (a driver call each) only then.
static int cached_tmap(bfk_index* idx, int slot, const void* base, int64_t rows, int box_rows, const CUtensorMap** out) {
    bfk_index::TmapSlot& c = idx->tmap[slot];
    static_assert(sizeof(CUtensorMap) <= sizeof(c.map), "tensor map cache slot too small");
    CUtensorMap* m = reinterpret_cast<CUtensorMap*>(c.map);
    if (c.base != base || c.rows != rows || c.box_rows != box_rows) {
        c.base = nullptr;
        BFK_TRY(make_tmap(m, base, rows, idx->ld16, box_rows));
        c.base = base;
        c.rows = rows;
        c.box_rows = box_rows;
    }
    *out = m;
    return 0;
}

struct TcPlan {
    int KB, stages, QT, S, KP, NS, G, J, grid, a_resident, pend_global, csize;
    int64_t nq_pad, tiles_total;
    size_t smem;
};

// pair_mode: -1 = default policy, 0 = single CTAs, 1 = CTA pairs when the number of query tiles is even
static int tc_plan(const bfk_index* idx, int64_t nq, int k, TcPlan* pl, int pair_mode = -1) {
    pl->KB = idx->ld16 / 64;
    // measured: moving the lists to global memory to win a 4th stage made flushes ~10x slower and the extra
    // stage bought nothing (the multi-tile regime is bound by L2 -> SM bandwidth, not by ring depth)
    pl->pend_global = 0;
    pl->QT = (int)((nq + TC_BM - 1) / TC_BM);
    // CTA pairs (needs an even number of query tiles): tcgen05.mma.cta_group::2 with M = 256 -- the pair
    // shares one corpus stream and each CTA stages half of every corpus tile, which halves the L2 -> SM bytes
    // per MMA (single CTAs run at ~87 % of the chip's L2 throughput cap at nq = 4096) and doubles the ring
    // depth in time.  Measured once the MMA issue loop stopped being the bottleneck: nq=1024 0.80 vs 0.77,
    // nq=2048 0.84 vs 0.81, nq=4096 0.76 vs 0.72 of the burst bf16 peak.  BFK_TC_PAIR=0 turns it off.
    static const int pair_env = [] {
        const char* e = getenv("BFK_TC_PAIR");
        return e ? (atoi(e) != 0 ? 1 : 0) : -1;
    }();
    if (pair_mode < 0) pair_mode = pair_env < 0 ? 1 : pair_env;
    pl->csize = (pair_mode == 1 && pl->QT % 2 == 0) ? 2 : 1;
    const size_t b_block = pl->csize == 2 ? TC_B_BLOCK / 2 : TC_B_BLOCK;
    // the 128-query tile stays resident in shared memory when it leaves room for >= 3 corpus stages;
    // otherwise (d > 384) its K-blocks stream through the ring next to the corpus K-blocks
    const size_t tail = (size_t)(pl->pend_global ? 0 : TC_PEND_BYTES) + TC_TAIL_BYTES;
    const size_t res_fixed = (size_t)pl->KB * TC_A_BLOCK + tail;
    pl->a_resident = (res_fixed + 3 * (size_t)TC_B_BLOCK <= TC_SMEM_LIMIT) ? 1 : 0;
    const size_t fixed = pl->a_resident ? res_fixed : tail;
    const size_t stage = pl->a_resident ? b_block : b_block + TC_A_BLOCK;
    if (fixed + 2 * stage > TC_SMEM_LIMIT) return 1;
    pl->stages = (int)std::min<size_t>(TC_MAX_STAGES, (TC_SMEM_LIMIT - fixed) / stage);
    pl->smem = fixed + (size_t)pl->stages * stage;
    pl->nq_pad = (int64_t)pl->QT * TC_BM;
    pl->tiles_total = (idx->ntotal + TC_BN - 1) / TC_BN;
    pl->S = (int)std::max<int64_t>(1, std::min<int64_t>(idx->num_sms / pl->QT, pl->tiles_total));
    pl->grid = pl->QT * pl->S;
    // KP list slots per (query, thread-stream).  The certificate is global: the streams' lists together
    // must hold the k best rows plus the error band, so KP may be smaller than k when there are many
    // streams; few streams get the longer lists.
    pl->KP = (k <= 12 && pl->S >= 8) ? 16 : 32;
    const int TS = 2 * pl->S;  // thread-streams per query
    if (k > 24 && (int64_t)TS * pl->KP < 2 * (int64_t)k) return 1;  // lists too short to certify: exact path
    // NS >= k published slots bound the global k-th best key; G groups x J ranks
    // (the fewer slots, the tighter the bound: max over NS distinct rows bounds the NS-th best)
    int NS = 16;
    while (NS < k + k / 4) NS <<= 1;
    int G = 1;
    while (G * 2 <= std::min(16, TS) && G * 2 <= NS) G *= 2;
    if (NS / G > pl->KP) G = NS / pl->KP;  // ranks per thread cannot exceed its list (extra groups stay unset)
    pl->NS = NS;
    pl->G = G;
    pl->J = NS / G;
    return 0;
}

// Queries per pass.  A pass runs QT query tiles x S = floor(SMs / QT) corpus streams, so it leaves SMs idle
// when SMs / QT is far from integral (QT = 32: 128 of 148).  Splitting the batch into passes of c tiles costs
// one more corpus read from HBM per pass (cheap next to the contraction once c >= 2) and a fixed launch /
// fill / drain overhead, and pays back when c * floor(SMs / c) covers the chip better: nq = 4096 runs as
// 2 x 2048 (144 SMs).  Cost model in units of one CTA's tile time: a pass walks ceil(T / S) tiles per CTA, is
// never faster than the HBM read of the corpus (~ T / 75 at d = 384), plus ~16 tiles of overhead.
static int64_t tc_chunk_queries(const bfk_index* idx, int64_t nq) {
    const int sms = idx->num_sms;
    const int64_t QT = (nq + TC_BM - 1) / TC_BM;
    const int64_t T = (idx->ntotal + TC_BN - 1) / TC_BN;
    auto pass_cost = [&](int64_t c) {
        const int64_t S = std::max<int64_t>(1, std::min<int64_t>(sms / c, T));
        // fewer than 8 streams per query tile run the long lists (KP = 32: more registers, costlier inserts):
        // measured 2410 us for nq = 4096 as 3072 + 1024 against 2335 us as 2 x 2048
        const double long_lists = S < 8 ? 1.05 : 1.0;
        // every pass reads the corpus from HBM once more.  That is never the bottleneck of a tensor-bound pass, but
        // it is not free either (L2 misses in front of the ring, and power on a power-capped part): config 3 as
        // 8 passes of 512 queries (148 SMs busy) measured 57.3 ms, as 2 passes of 2048 (144 SMs) 52.3 ms.  A third of
        // the corpus read time charged on top of the compute time reproduces both that and the 1M x 384 measurements.
        return ((double)std::max<int64_t>((T + S - 1) / S, T / 75) + 16.0) * long_lists + 0.3 * (double)(T / 75);
    };
    static const int forced = [] {  // experiments: BFK_TC_CHUNK_TILES=c forces passes of c query tiles
        const char* e = getenv("BFK_TC_CHUNK_TILES");
        return e ? atoi(e) : 0;
    }();
    if (forced > 0) return std::min<int64_t>(forced, std::min<int64_t>(QT, sms)) * TC_BM;
    int64_t best_c = std::min<int64_t>(QT, sms);
    double best = -1.0;
    for (int64_t c = std::min<int64_t>(QT, sms); c >= 1; c--) {
        const int64_t full = QT / c, rest = QT - full * c;
        // even tile counts run as CTA pairs (measured ~4 % faster)
        const double cost = (double)full * pass_cost(c) * ((c & 1) ? 1.04 : 1.0) + (rest ? pass_cost(rest) : 0.0);
        if (best < 0 || cost < best * 0.995) {  // ties go to the larger pass (the model tracks measurements within ~2 %)
            best = cost;
            best_c = c;
        }
    }
    return best_c * TC_BM;
}

int tensor_path_supported(const bfk_index* idx, int64_t nq, int k) {
    if (k > 128 || nq <= 0 || idx->ntotal <= 0) return 0;
    TcPlan pl;
    if (tc_plan(idx, std::min<int64_t>(nq, tc_chunk_queries(idx, nq)), k, &pl)) return 0;
    return get_encode_fn() != nullptr;
}

struct TcWs {
    __nv_bfloat16* q16;
    float* qn;
    unsigned int* slots;
    float* qmargin;
    uint64_t* out_cand;
    float* out_thr;
    int* flags;      // [0] = count, [1..] = list
    float* fb_q;     // [8][ldq] compacted fallback queries
    float* fb_D;     // [8][k]
    int64_t* fb_I;   // [8][k]
    uint64_t* fb_part;
    unsigned long long* dbg;
    uint64_t* pend_g;
    // second-chance pass over the uncertified queries only (fewer queries -> more streams -> more list room)
    int* r_list;     // [nq/2 + 1] query indices (copy of the first pass's flag list)
    float* r_q;      // [nq/2][ldq]
    float* r_D;      // [nq/2][k]
    int64_t* r_I;    // [nq/2][k]
    size_t bytes;
};

// sizes depend on the query count only (list space is bounded by the grid), so a workspace carved for
// nqc queries also fits every smaller second-chance pass
static void tc_carve(const bfk_index* idx, int64_t nqc, int k, int fb_nparts, char* base, TcWs* w) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        off = (size_t)round_up((int64_t)off, 256);
        char* p = base ? base + off : nullptr;
        off += bytes;
        return p;
    };
    const int64_t nq_pad = round_up(nqc, TC_BM);
    int NS = 32;
    while (NS < k + k / 4) NS <<= 1;
    const size_t max_lists = (size_t)idx->num_sms * TC_BM * 2;  // thread-streams per launch, at most
    w->q16 = (__nv_bfloat16*)take((size_t)nq_pad * idx->ld16 * 2);
    w->qn = (float*)take((size_t)nq_pad * 4);
    w->slots = (unsigned int*)take((size_t)nq_pad * NS * 4);
    w->qmargin = (float*)take((size_t)nq_pad * 4);
    w->out_cand = (uint64_t*)take(max_lists * 32 * 8);
    w->out_thr = (float*)take(max_lists * 4);
    w->flags = (int*)take((size_t)(nqc + 2) * 4) + 1;  // [-1] ticket, [0] count, [1..] list
    w->fb_q = (float*)take((size_t)8 * idx->ld32 * 4);
    w->fb_D = (float*)take((size_t)8 * k * 4);
    w->fb_I = (int64_t*)take((size_t)8 * k * 8);
    w->fb_part = (uint64_t*)take((size_t)8 * fb_nparts * k * 8);
    w->dbg = (unsigned long long*)take((size_t)idx->num_sms * 8 * 8 + 8 * 64 * 8 + 16 * 8 + (size_t)nq_pad * 4 * 8);  // scan counters, timeline, K5 phases, K5 per query
    w->pend_g = (uint64_t*)take((size_t)idx->num_sms * TC_BUFCAP * TC_EPI_THREADS * 8);
    const int64_t nr = nqc / 2 + 1;
    w->r_list = (int*)take((size_t)(nr + 1) * 4);
    w->r_q = (float*)take((size_t)nr * idx->ld32 * 4);
    w->r_D = (float*)take((size_t)nr * k * 4);
    w->r_I = (int64_t*)take((size_t)nr * k * 8);
    w->bytes = off + 256;
}

size_t tensor_workspace_bytes(const bfk_index* idx, int64_t nq, int k) {
    const int64_t chunk = std::min<int64_t>(nq, tc_chunk_queries(idx, nq));
    size_t sb = 0, bytes = 0;
    int nparts = 0;
    if (scan_exact_workspace(idx, 8, k, &sb, &nparts)) return 0;
    // the full passes and the shorter last one are planned separately: size for the larger of the two
    const int64_t sizes[2] = {chunk, nq % chunk};
    for (int64_t nqc : sizes) {
        if (nqc <= 0) continue;
        TcPlan pl;
        if (tc_plan(idx, nqc, k, &pl)) return 0;
        TcWs w;
        tc_carve(idx, nqc, k, nparts, nullptr, &w);
        bytes = std::max(bytes, w.bytes);
    }
    return bytes;
}

// co-resident clusters of `csize` CTAs for this kernel and shared-memory size (cluster size 4 strands SMs
// in GPCs whose SM count is not a multiple of 4: 132 of 148 on B200)
template <int KP>
static int max_cluster_ctas(int csize, size_t smem) {
    auto kern = scan_tc_kernel<KP, true, false>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_LIMIT);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(148 * 4);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)csize;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n * csize;
}

// BFK_TC_PDL=0 launches the three kernels of a pass strictly one after the other
static bool pdl_enabled() {
    static const bool on = [] {
        const char* e = getenv("BFK_TC_PDL");
        return e ? atoi(e) != 0 : true;
    }();
    return on;
}

template <int KP, bool PAIR, bool DBG>
static int launch_scan_tc(const CUtensorMap& tmQ, const CUtensorMap& tmX, const TcParams& p, const TcPlan& pl,
                          cudaStream_t st, bool pdl) {
    auto kern = scan_tc_kernel<KP, PAIR, DBG>;
    static bool attr_set[64] = {false};  // the attribute is per device
    int dev = 0;
    BFK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        BFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_LIMIT));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)pl.grid);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)pl.csize;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    // programmatic dependent launch behind prep_queries_kernel: the prologue (barriers, TMEM, descriptor
    // prefetch) overlaps its tail; the kernel executes griddepcontrol.wait before it reads prep's outputs
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 2 : 1;
    BFK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmQ, tmX, p));
    BFK_CUDA(cudaGetLastError());
    return 0;
}

// rows of (rD, rI) [n][k] -> rows list[i] of (D, I); and queries list[i] -> compact rows
__global__ void __launch_bounds__(256) gather_queries_kernel(const float* __restrict__ q, int ldq, const int* __restrict__ list,
                                                             int n, float* __restrict__ out) {
    const int64_t total = (int64_t)n * ldq;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(i / ldq), c = (int)(i - (int64_t)r * ldq);
        out[i] = q[(size_t)list[r] * ldq + c];
    }
}
__global__ void __launch_bounds__(256) scatter_results_kernel(const float* __restrict__ rD, const int64_t* __restrict__ rI,
                                                              const int* __restrict__ list, int n, int k,
                                                              float* __restrict__ D, int64_t* __restrict__ I) {
    const int64_t total = (int64_t)n * k;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(i / k), c = (int)(i - (int64_t)r * k);
        D[(size_t)list[r] * k + c] = rD[i];
        I[(size_t)list[r] * k + c] = rI[i];
    }
}

// one tensor pass: prep -> K3b scan -> K5 rerank, all enqueued on `st`; w.flags[0] = #uncertified,
// w.flags[1..] = their indices (relative to qc)
// q_src / ld_src: where prep reads the fp32 queries -- qc itself, or pinned host memory (then qc receives a copy);
// out_is_host: D_out / I_out are pinned host memory that K5 writes directly
static int tc_pass(bfk_index* idx, const float* qc, int ldq, int64_t nqc, int k, float* D_out, int64_t* I_out,
                   int64_t id_offset, const TcWs& w, const TcPlan& plan, cudaStream_t st, cudaEvent_t ev0, cudaEvent_t ev1,
                   int* launches, const float* q_src = nullptr, int ld_src = 0, int out_is_host = 0) {
    const bool l2 = idx->metric == BFK_METRIC_L2;
    TcPlan pl = plan;
    // every CTA of the persistent grid must be resident at once for the cluster size chosen; when the pairs do
    // not all fit (MIG slices, SMs held by another context) the same search runs with single CTAs
    if (pl.csize == 2) {
        struct Fit { int dev, kp; size_t smem; int fit; };
        static std::mutex fit_mu;
        static std::vector<Fit> fits;
        int fit = -1;
        {
            std::lock_guard<std::mutex> g(fit_mu);
            for (const Fit& f : fits)
                if (f.dev == idx->device && f.kp == pl.KP && f.smem == pl.smem) fit = f.fit;
            if (fit < 0) {
                fit = pl.KP == 16 ? max_cluster_ctas<16>(2, pl.smem) : max_cluster_ctas<32>(2, pl.smem);
                fits.push_back({idx->device, pl.KP, pl.smem, fit});
            }
        }
        if (fit < pl.grid) {
            TcPlan single;
            if (tc_plan(idx, nqc, k, &single, 0)) return fail("tensor path: shape not supported");
            pl = single;
        }
    }
    // eps = bound on |<q,x> - acc|, acc = what the tensor pipe returns for <bf16 q, scan copy of x>; put together per
    // query in prep_queries_kernel from measured rounding losses (see there) plus the accumulation term:
    //   the products of two bf16 numbers are exact in fp32; the pipe adds ld16 of them in fp32.  Assuming nothing
    //   better than truncation (round toward zero, one ulp = 2^-23 relative per addition, the weaker of the two
    //   rounding modes it could use), the sum is off by at most ld16 * 2^-23 * sum|q_i x_i| <= ld16 * 2^-23 ||q|| ||x||.
    //   That is 2^-14 at ld16 = 512 -- the constant used for every d before -- and 2^-12 at BFK_MAX_D = 2048, so it
    //   is kept as a floor and grows with the row length above it.
    const float c_acc = (float)std::max(idx->ld16, 512) / 8388608.f;
    unsigned long long* ts = nullptr;  // this pass's device-clock stamps (bfk_profile_begin), or none
    if (idx->prof_on && idx->prof_ring && idx->prof_used < idx->prof_cap) ts = idx->prof_ring + 4 * idx->prof_used++;
    // the fp32 re-score itself (ld16/32 terms per lane, then a 5-step butterfly) and the key's own fma round
    // with at most (ld16/32 + 8) half-ulps of a value that is at most 2 (||q||^2 + ||x||^2)
    const float c_abs = (float)(idx->ld16 / 32 + 8) * 2.f / 16777216.f;
    const bool q_remote = q_src != nullptr && q_src != qc;
    prep_queries_kernel<<<(unsigned)((pl.nq_pad + 7) / 8), 256, 0, st>>>(
        q_remote ? q_src : qc, q_remote ? ld_src : ldq, idx->d, nqc, pl.nq_pad, w.q16, idx->ld16, w.qn, idx->maxn, c_acc,
        c_abs, l2 ? 1 : 0, w.qmargin, w.slots, pl.NS, w.flags, ts, q_remote ? const_cast<float*>(qc) : nullptr, ldq);
    BFK_CUDA(cudaGetLastError());
    const CUtensorMap *tmQp, *tmXp;
    BFK_TRY(cached_tmap(idx, 1, w.q16, pl.nq_pad, TC_BM, &tmQp));
    BFK_TRY(cached_tmap(idx, 0, idx->x16, idx->ntotal, TC_BN / pl.csize, &tmXp));  // box = this CTA's slice
    const CUtensorMap& tmQ = *tmQp;
    const CUtensorMap& tmX = *tmXp;
    TcParams p;
    p.nrm = idx->sel_bias ? idx->sel_bias : idx->nrm;  // a selector search masks rows through the bias
    p.use_bias = (l2 || idx->sel_bias) ? 1 : 0;
    p.n = idx->ntotal;
    p.KB = pl.KB;
    p.QT = pl.QT;
    p.S = pl.S;
    p.tiles_total = pl.tiles_total;
    p.smem_bytes = (int)pl.smem;
    p.is_l2 = l2 ? 1 : 0;
    p.nq = (int)nqc;
    p.slots = w.slots;
    p.NS = pl.NS;
    p.G = pl.G;
    p.J = pl.J;
    p.a_resident = pl.a_resident;
    p.qmargin = w.qmargin;
    p.out_cand = w.out_cand;
    p.out_thr = w.out_thr;
    p.pend_g = pl.pend_global ? w.pend_g : nullptr;
    p.csize = pl.csize;
    p.prepass = (pl.J == 1 && pl.G >= k && 2 * pl.S >= pl.G) ? 1 : 0;
    static const bool prepass_on = [] {
        const char* e = getenv("BFK_TC_PREPASS");
        return e ? atoi(e) != 0 : true;
    }();
    if (!prepass_on) p.prepass = 0;
    p.tstamp = ts;
    static const bool dbg_on = getenv("BFK_TC_DEBUG") != nullptr;
    static const bool parked = getenv("BFK_TC_PARKED") != nullptr;
    p.dbg = dbg_on ? w.dbg : nullptr;
    p.parked = parked ? 1 : 0;
    if (dbg_on) BFK_CUDA(cudaMemsetAsync(w.dbg, 0, (size_t)pl.grid * 64 + 8 * 64 * 8, st));
    if (ev0) cudaEventRecord(ev0, st);
    // the instrumented build of the kernel (wait-cycle counters, timeline marks, parked waits) is a
    // separate instantiation: the single MMA-issuing thread paces the tensor pipe, and even predicated-off
    // diagnostics in its loop cost issue slots
    const bool dbgk = dbg_on || parked;
    // an event recorded between two kernels keeps them from overlapping: timed / instrumented passes run serially
    const bool pdl = pdl_enabled() && !ev0 && !ev1 && !dbg_on;
    if (pl.KP == 16) {
        if (pl.csize == 2 && dbgk) BFK_TRY((launch_scan_tc<16, true, true>(tmQ, tmX, p, pl, st, pdl)));
        else if (pl.csize == 2) BFK_TRY((launch_scan_tc<16, true, false>(tmQ, tmX, p, pl, st, pdl)));
        else if (dbgk) BFK_TRY((launch_scan_tc<16, false, true>(tmQ, tmX, p, pl, st, pdl)));
        else BFK_TRY((launch_scan_tc<16, false, false>(tmQ, tmX, p, pl, st, pdl)));
    } else {
        if (pl.csize == 2 && dbgk) BFK_TRY((launch_scan_tc<32, true, true>(tmQ, tmX, p, pl, st, pdl)));
        else if (pl.csize == 2) BFK_TRY((launch_scan_tc<32, true, false>(tmQ, tmX, p, pl, st, pdl)));
        else if (dbgk) BFK_TRY((launch_scan_tc<32, false, true>(tmQ, tmX, p, pl, st, pdl)));
        else BFK_TRY((launch_scan_tc<32, false, false>(tmQ, tmX, p, pl, st, pdl)));
    }
    if (ev1) cudaEventRecord(ev1, st);
    if (p.dbg) {  // diagnostics only: per-role wait cycles averaged over the CTAs
        std::vector<unsigned long long> h((size_t)pl.grid * 8 + 8 * 64);
        BFK_CUDA(cudaMemcpyAsync(h.data(), w.dbg, h.size() * 8, cudaMemcpyDeviceToHost, st));
        BFK_CUDA(cudaStreamSynchronize(st));
        double a[8] = {0};
        for (int c = 0; c < pl.grid; c++)
            for (int j = 0; j < 8; j++) a[j] += (double)h[(size_t)c * 8 + j] / pl.grid;
        fprintf(stderr,
                "[bfk tc] nq=%lld grid=%d QT=%d S=%d KP=%d NS=%d G=%d J=%d tiles/cta=%.1f | cycles/cta: life=%.0f "
                "tma_wait_empty=%.0f mma_wait_full=%.0f mma_wait_tmem=%.0f epi_wait_acc=%.0f epi_flush=%.0f | "
                "appended/thread=%.0f flushes=%.1f\n",
                (long long)nqc, pl.grid, pl.QT, pl.S, pl.KP, pl.NS, pl.G, pl.J, (double)pl.tiles_total / pl.S, a[7], a[0],
                a[1], a[2], a[3], a[4], a[5], a[6]);
        if (getenv("BFK_TC_TIMELINE")) {
            const unsigned long long* tl = h.data() + (size_t)pl.grid * 8;
            fprintf(stderr, "[bfk tc timeline, CTA 0, cycles] tile tma_first tma_last mma_acq mma_kb0 mma_kbL epi_start epi_done\n");
            for (int t = 8; t < 24; t++)
                fprintf(stderr, "  %2d %8llu %8llu %8llu %8llu %8llu %8llu %8llu\n", t, tl[0 * 64 + t], tl[1 * 64 + t],
                        tl[2 * 64 + t], tl[3 * 64 + t], tl[4 * 64 + t], tl[5 * 64 + t], tl[6 * 64 + t]);
            fprintf(stderr, "[tile 12, MMA thread] kb wait_done fenced mmas_issued committed\n");
            for (int kb = 0; kb < std::min(pl.KB, 8); kb++)
                fprintf(stderr, "  %d %8llu %8llu %8llu %8llu\n", kb, tl[7 * 64 + kb * 4], tl[7 * 64 + kb * 4 + 1],
                        tl[7 * 64 + kb * 4 + 2], tl[7 * 64 + kb * 4 + 3]);
        }
    }
    RerankParams r;
    r.out_cand = w.out_cand;
    r.out_thr = w.out_thr;
    r.qmargin = w.qmargin;
    r.q = qc;
    r.ldq = ldq;
    r.S = 2 * pl.S;  // thread-streams
    r.KP = pl.KP;
    r.k = k;
    r.is_l2 = l2 ? 1 : 0;
    r.storage = idx->storage;
    r.x32 = idx->x32;
    r.ld32 = idx->ld32;
    r.x16 = idx->x16;
    r.ld16 = idx->ld16;
    r.d = idx->d;
    r.id_offset = id_offset;
    r.D = D_out;
    r.I = I_out;
    r.nflag = w.flags;
    r.flag_list = w.flags + 1;
    r.out_is_host = out_is_host;
    r.h_nflag = idx->h_flag_dev;
    r.seq = ++idx->flag_seq;
    r.dbg = p.dbg ? w.dbg + (size_t)pl.grid * 8 + 8 * 64 : nullptr;
    r.tstamp = ts;
    {
        const bool small = (int64_t)r.S * r.KP <= 1024 && k <= 32;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)nqc);
        cfg.blockDim = dim3(small ? 128 : 512);
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = pdl ? 1 : 0;
        if (small) BFK_CUDA(cudaLaunchKernelEx(&cfg, rerank_kernel<128>, r));
        else BFK_CUDA(cudaLaunchKernelEx(&cfg, rerank_kernel<512>, r));
    }
    BFK_CUDA(cudaGetLastError());
    if (r.dbg) {
        unsigned long long h[16];
        BFK_CUDA(cudaMemcpyAsync(h, r.dbg, sizeof(h), cudaMemcpyDeviceToHost, st));
        std::vector<unsigned long long> pq((size_t)nqc * 4);
        BFK_CUDA(cudaMemcpyAsync(pq.data(), r.dbg + 16, pq.size() * 8, cudaMemcpyDeviceToHost, st));
        BFK_CUDA(cudaStreamSynchronize(st));
        {
            std::vector<unsigned long long> life, rows, surv, start;
            for (int64_t i = 0; i < nqc; i++) {
                life.push_back(pq[4 * i]);
                rows.push_back(pq[4 * i + 1]);
                surv.push_back(pq[4 * i + 2]);
                start.push_back(pq[4 * i + 3]);
            }
            auto srt = [](std::vector<unsigned long long>& v) { std::sort(v.begin(), v.end()); };
            srt(life); srt(rows); srt(surv); srt(start);
            const size_t m = life.size() / 2, l = life.size() - 1;
            fprintf(stderr, "[bfk rerank, all %lld CTAs] lifetime cycles min/med/max %llu/%llu/%llu | rows re-scored %llu/%llu/%llu | "
                            "filter survivors %llu/%llu/%llu | start spread %llu ns | margin of query 0 (key space) %.6f\n",
                    (long long)nqc, life[0], life[m], life[l], rows[0], rows[m], rows[l], surv[0], surv[m], surv[l],
                    start[l] - start[0], (double)([&] { const uint32_t u = (uint32_t)h[10]; float f; memcpy(&f, &u, 4); return f; }()));
        }
        fprintf(stderr, "[bfk rerank, CTA 0, cycles] tmin=%llu filtered=%llu sorted=%llu cut=%llu rescored=%llu sorted2=%llu | survivors=%llu R=%llu\n",
                h[0], h[1], h[2], h[3], h[4], h[5], h[8], h[9]);
    }
    *launches += 3;
    return 0;
}

// everything needed to settle the uncertified queries of one tensor pass after the fact
struct TcPending {
    const float* qc;
    int ldq;
    int64_t nqc;
    int k;
    float* Dc;
    int64_t* Ic;
    int64_t id_offset;
    TcWs w;
    TcPlan pl;
    int fb_nparts;
    int out_is_host = 0;  // Dc / Ic are pinned host memory
};

// Reads the pass's flag list (one stream sync) and settles what was not certified: a second tensor pass
// over those queries alone, then the exact scan for what is still open.  Returns how many result rows
// were rewritten through *n_rewritten.
// Waits until the last tensor pass enqueued on `st` has published its certificate count.  K5's last CTA writes the
// count and then the pass's sequence number into pinned host memory; polling that word costs a PCIe write's
// latency, where cudaStreamSynchronize also waits for the kernel to be retired and for the driver to notice.
// Passes longer than a few hundred microseconds (and any error) fall back to the blocking wait.
int wait_flag_seq(bfk_index* idx, cudaStream_t st) {
    static const bool spin_on = [] {
        const char* e = getenv("BFK_TC_SPIN");
        return e ? atoi(e) != 0 : true;
    }();
    const int want = idx->flag_seq;
    volatile int* seq = idx->h_flag + 1;
    if (spin_on) {
        for (int round = 0; round < 64 && *seq != want; round++) {
            for (int i = 0; i < 400 && *seq != want; i++) {
#if defined(__x86_64__)
                __builtin_ia32_pause();
#endif
            }
            if (*seq != want && cudaStreamQuery(st) != cudaErrorNotReady) break;  // finished, or failed
        }
        cudaGetLastError();
        if (*seq == want) return 0;
    }
    BFK_CUDA(cudaStreamSynchronize(st));
    return 0;
}

static int tc_resolve(bfk_index* idx, const TcPending& c, cudaStream_t st, int64_t* n_retry, int64_t* n_fallback,
                      int64_t* n_rewritten, int* launches) {
    const TcWs& w = c.w;
    BFK_TRY(wait_flag_seq(idx, st));
    int nflag = *idx->h_flag;
    if (n_rewritten) *n_rewritten += nflag;
    if (nflag == 0) return 0;
    std::vector<int> list((size_t)nflag);
    BFK_CUDA(cudaMemcpyAsync(list.data(), w.flags + 1, sizeof(int) * (size_t)nflag, cudaMemcpyDeviceToHost, st));
    BFK_CUDA(cudaStreamSynchronize(st));

    // second chance: the same tensor pass over the uncertified queries alone.  Fewer queries means
    // more corpus streams per query, hence more list room and a better chance to certify -- for the
    // price of one HBM-bound corpus pass instead of nflag/8 exact passes.
    TcPlan pl2;
    if (2 * (int64_t)nflag <= c.nqc && !tc_plan(idx, nflag, c.k, &pl2) && pl2.S >= 2 * c.pl.S) {
        BFK_CUDA(cudaMemcpyAsync(w.r_list, w.flags + 1, sizeof(int) * (size_t)nflag, cudaMemcpyDeviceToDevice, st));
        gather_queries_kernel<<<(unsigned)std::min<int64_t>(((int64_t)nflag * c.ldq + 255) / 256, 1184), 256, 0, st>>>(
            c.qc, c.ldq, w.r_list, nflag, w.r_q);
        BFK_CUDA(cudaGetLastError());
        BFK_TRY(tc_pass(idx, w.r_q, c.ldq, nflag, c.k, w.r_D, w.r_I, c.id_offset, w, pl2, st, nullptr, nullptr, launches));
        scatter_results_kernel<<<(unsigned)std::min<int64_t>(((int64_t)nflag * c.k + 255) / 256, 1184), 256, 0, st>>>(
            w.r_D, w.r_I, w.r_list, nflag, c.k, c.Dc, c.Ic);
        BFK_CUDA(cudaGetLastError());
        *launches += 2;
        BFK_CUDA(cudaStreamSynchronize(st));  // the scatter above has to land, too
        const int nflag2 = *idx->h_flag;
        std::vector<int> list2((size_t)nflag2);
        if (nflag2 > 0) {
            BFK_CUDA(cudaMemcpyAsync(list2.data(), w.flags + 1, sizeof(int) * (size_t)nflag2, cudaMemcpyDeviceToHost, st));
            BFK_CUDA(cudaStreamSynchronize(st));
        }
        *n_retry += nflag;
        std::vector<int> still;
        still.reserve((size_t)nflag2);
        for (int j : list2) still.push_back(list[(size_t)j]);
        list.swap(still);
        nflag = nflag2;
    }

    // ---- last resort: exact re-run on the GPU (K3a + K4), 8 queries per corpus pass ----
    std::sort(list.begin(), list.end());
    for (int g0 = 0; g0 < nflag; g0 += 8) {
        const int g = std::min(8, nflag - g0);
        for (int j = 0; j < g; j++)
            BFK_CUDA(cudaMemcpyAsync(w.fb_q + (size_t)j * c.ldq, c.qc + (size_t)list[g0 + j] * c.ldq, sizeof(float) * c.ldq,
                                     cudaMemcpyDeviceToDevice, st));
        BFK_TRY(launch_scan_exact(idx, w.fb_q, c.ldq, g, c.k, w.fb_part, c.fb_nparts, st, idx->sel));
        BFK_TRY(launch_select(w.fb_part, (int64_t)c.fb_nparts * c.k, g, c.k, idx->metric, c.id_offset, w.fb_D, w.fb_I, c.k, st));
        *launches += 2;
        for (int j = 0; j < g; j++) {
            const size_t row = (size_t)list[g0 + j] * c.k;
            // (the destination may be pinned host memory: cudaMemcpyDefault)
            BFK_CUDA(cudaMemcpyAsync(c.Dc + row, w.fb_D + (size_t)j * c.k, sizeof(float) * c.k, cudaMemcpyDefault, st));
            BFK_CUDA(cudaMemcpyAsync(c.Ic + row, w.fb_I + (size_t)j * c.k, sizeof(int64_t) * c.k, cudaMemcpyDefault, st));
        }
    }
    *n_fallback += nflag;
    // rewrites into pinned host memory are copies the caller cannot see coming: they have landed on return
    if (c.out_is_host) BFK_CUDA(cudaStreamSynchronize(st));
    return 0;
}

// one deferred pass per index at most; settled by search_tensor_finish (or by the next search)
static std::mutex g_pending_mu;
static std::vector<std::pair<bfk_index*, TcPending>> g_pending;

int search_tensor_finish(bfk_index* idx, cudaStream_t st, int64_t* n_rewritten) {
    TcPending c;
    bool have = false;
    {
        std::lock_guard<std::mutex> lk(g_pending_mu);
        for (size_t i = 0; i < g_pending.size(); i++)
            if (g_pending[i].first == idx) {
                c = g_pending[i].second;
                g_pending.erase(g_pending.begin() + (long)i);
                have = true;
                break;
            }
    }
    if (n_rewritten) *n_rewritten = 0;
    if (!have) return 0;
    int launches = 0;
    int64_t n_retry = 0, n_fallback = 0;
    return tc_resolve(idx, c, st, &n_retry, &n_fallback, n_rewritten, &launches);
}

// device address of the uncertified-query count of the pass a deferred search left pending (or nullptr)
const int* search_tensor_pending_nflag(bfk_index* idx) {
    std::lock_guard<std::mutex> lk(g_pending_mu);
    for (size_t i = 0; i < g_pending.size(); i++)
        if (g_pending[i].first == idx) return g_pending[i].second.w.flags;
    return nullptr;
}

void search_tensor_forget(bfk_index* idx) {
    std::lock_guard<std::mutex> lk(g_pending_mu);
    for (size_t i = 0; i < g_pending.size(); i++)
        if (g_pending[i].first == idx) {
            g_pending.erase(g_pending.begin() + (long)i);
            break;
        }
}

// defer != 0 (single chunk only): enqueue the pass and return without looking at the flags; the caller
// must settle with search_tensor_finish before trusting the results.
int search_tensor(bfk_index* idx, const float* q_dev, int ldq, int64_t nq, int k, float* D_dev, int64_t* I_dev,
                  int64_t id_offset, char* ws_tail, size_t ws_tail_bytes, cudaStream_t st, bfk_search_stats* stats,
                  cudaEvent_t ev0, cudaEvent_t ev1, int defer, const float* q_host, int out_is_host) {
    const int64_t chunk = tc_chunk_queries(idx, nq);
    int launches = 0;
    int64_t n_fallback = 0, n_retry = 0;
    int fb_nparts = 0;
    size_t sb = 0;
    BFK_TRY(scan_exact_workspace(idx, 8, k, &sb, &fb_nparts));
    for (int64_t q0 = 0; q0 < nq; q0 += chunk) {
        const int64_t nqc = std::min(chunk, nq - q0);
        TcPending c;
        if (tc_plan(idx, nqc, k, &c.pl)) return fail("tensor path: shape not supported");
        tc_carve(idx, nqc, k, fb_nparts, ws_tail, &c.w);
        if (c.w.bytes > ws_tail_bytes) return fail("internal: tensor-path workspace under-sized");
        c.qc = q_dev + (size_t)q0 * ldq;
        c.ldq = ldq;
        c.nqc = nqc;
        c.k = k;
        c.Dc = D_dev + (size_t)q0 * k;
        c.Ic = I_dev + (size_t)q0 * k;
        c.id_offset = id_offset;
        c.fb_nparts = fb_nparts;
        c.out_is_host = out_is_host;
        BFK_TRY(tc_pass(idx, c.qc, ldq, nqc, k, c.Dc, c.Ic, id_offset, c.w, c.pl, st, q0 == 0 ? ev0 : nullptr,
                        q0 + chunk >= nq ? ev1 : nullptr, &launches, q_host ? q_host + (size_t)q0 * idx->d : nullptr,
                        idx->d, out_is_host));
        if (defer && nq <= chunk) {
            std::lock_guard<std::mutex> lk(g_pending_mu);
            g_pending.emplace_back(idx, c);
            break;
        }
        BFK_TRY(tc_resolve(idx, c, st, &n_retry, &n_fallback, nullptr, &launches));
    }
    if (stats) {
        stats->n_launches = launches;
        stats->n_fallback = n_fallback;
        stats->n_candidates = n_retry;
        stats->algo_used = BFK_ALGO_TENSOR;
    }
    return 0;
}

}  // namespace bfk

// K3a `scan_exact`: exact fp32 flat scan for up to 8 queries per corpus pass.
//
// Stands in for faiss IndexFlat::search's sequential path (knn_L2sqr ->
// exhaustive_L2sqr_seq / fvec_L2sqr, and the inner-product twin), which is what the
// reference's only search call runs: llm-qa/main.py:101,117 issues
// index.search(float32[1, 384], k=3) through LangChain.  Like that path it forms the
// direct difference sum((x - y)^2) in fp32, so duplicates give exactly 0.
//
// Shape of the kernel (HBM-bound: 4*ld32 bytes per row, one pass per <= 8 queries):
//   * persistent grid, one CTA per SM, 9 warps: warp 8 is the copy producer, warps 0..7
//     are consumers;
//   * corpus rows are contiguous, so a tile of TR = 8*R rows is ONE `cp.async.bulk`
//     (TMA 1-D bulk copy) global -> shared, completion on an mbarrier; `stages` tiles
//     are in flight per SM, which is what keeps HBM busy without register staging;
//   * each consumer warp owns R rows of the tile, reads them with conflict-free
//     LDS.128, keeps R*NQ partial sums per lane and reduces them with a transposing
//     butterfly (V-1 + 5-log2(V) shuffles instead of 5*V);
//   * selection: per query a shared candidate buffer of composite keys
//     (ordered distance << 32 | row id) with a threshold = current k-th best; when the
//     buffer can no longer take a tile's worth of inserts the consumers bitonic-sort
//     it and keep the best k.  Order is total on (distance, id): results do not depend
//     on scheduling.
//   * each CTA writes its sorted top-k per query to part[query][cta][k]; K4
//     (bfk_select.cu) merges the CTAs' lists.
#include <cfloat>

#include "bfk_common.cuh"
#include "bfk_ptx.cuh"

namespace bfk {

constexpr int SE_CWARPS = 8;                       // consumer warps
constexpr int SE_CTHREADS = SE_CWARPS * 32;        // 256
constexpr int SE_THREADS = SE_CTHREADS + 32;       // + producer warp
constexpr int SE_SMEM_BUDGET = 220 * 1024;
constexpr int SE_MAX_STAGES = 6;

// consumers-only barrier (named barrier 1); the producer warp never joins it
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(SE_CTHREADS) : "memory"); }

// in-place ascending bitonic sort of a[0..n) (n power of two) by the consumer threads
__device__ void consumer_bitonic_sort(uint64_t* a, int n, int tid) {
    for (int k2 = 2; k2 <= n; k2 <<= 1) {
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < (n >> 1); i += SE_CTHREADS) {
                int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
                int hi = lo + j;
                bool asc = (lo & k2) == 0;
                uint64_t x = a[lo], y = a[hi];
                if ((x > y) == asc) {
                    a[lo] = y;
                    a[hi] = x;
                }
            }
            consumer_sync();
        }
    }
}

// the same for short arrays (n <= 256, a power of two): every element's rank is counted -- n*n independent compares
// over the 256 consumer threads, three barriers in all -- instead of a network of up to 36 barrier-separated steps.
// Equal composites (only COMP_EMPTY padding repeats) are ordered by position, so ranks are distinct.
__device__ void consumer_sort(uint64_t* a, int n, int tid) {
    if (n > SE_CTHREADS) {
        consumer_bitonic_sort(a, n, tid);
        return;
    }
    const int P = SE_CTHREADS / n;  // threads per element: neighbours in one warp (P <= 8 as n >= 32)
    const int e = tid / P, part = tid - e * P;
    const uint64_t v = a[e];
    int cnt = 0;
    for (int j = part; j < n; j += P) {
        const uint64_t o = a[j];
        cnt += (o < v || (o == v && j < e)) ? 1 : 0;
    }
    for (int o = P >> 1; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    consumer_sync();
    if (part == 0) a[cnt] = v;
    consumer_sync();
}

template <typename RowT>
struct RowVec;
template <>
struct RowVec<float> {
    static constexpr int VEC = 4;
    __device__ static __forceinline__ void load(const unsigned char* p, float (&v)[4]) {
        float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
};
template <>
struct RowVec<__nv_bfloat16> {
    static constexpr int VEC = 8;
    __device__ static __forceinline__ void load(const unsigned char* p, float (&v)[8]) {
        uint4 t = *reinterpret_cast<const uint4*>(p);
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
};

struct ScanExactParams {
    const void* xb;      // corpus rows
    int64_t n;           // rows
    int ld;              // elements per row (multiple of the 16-byte vector)
    const float* q;      // [nq_real, ldq] fp32, zero padded to ldq >= ld
    int ldq;
    int nq_real;
    int k;
    int cap;             // candidate slots per query (power of two, >= k + tile_rows)
    int stages;
    uint64_t* part;      // [nq_real][nparts][k]
    int nparts;
    const unsigned char* sel;  // optional id selector bitmap (bit i <=> row i admitted), else null
    // K4 folded into this kernel (small searches: the second launch and the gap before it cost more than the merge):
    // the last CTA to publish its lists merges them and writes (D, I) itself
    int fuse;
    int metric;
    int merge_n;               // power of two >= nparts * k (and >= k): the sort window
    int64_t id_offset;
    float* D;                  // [nq_real][out_ld]
    int64_t* I;
    int64_t out_ld;
    int* ticket;               // device word, 0 between searches
    int* h_seq;                // optional pinned host word: the sequence number below is written there at the end
    int seq;
};

template <int NQ, int R, bool IS_L2, typename RowT>
__global__ void __launch_bounds__(SE_THREADS, 1) scan_exact_kernel(const ScanExactParams p) {
    constexpr int VEC = RowVec<RowT>::VEC;
    constexpr int TR = SE_CWARPS * R;  // rows per tile
    constexpr int V = R * NQ;          // partial sums per lane
    static_assert((V & (V - 1)) == 0 && V <= 32, "V must be a power of two <= 32");

    extern __shared__ __align__(128) unsigned char smem[];
    const int row_bytes = p.ld * (int)sizeof(RowT);
    const int tile_bytes = TR * row_bytes;
    unsigned char* tiles = smem;                                            // [stages][tile_bytes]
    float* q_s = reinterpret_cast<float*>(tiles + (size_t)p.stages * tile_bytes);  // [NQ][ld]
    uint64_t* cand = reinterpret_cast<uint64_t*>(q_s + NQ * p.ld);          // [NQ][cap]
    uint64_t* thr_s = cand + (size_t)NQ * p.cap;                            // [NQ]
    uint64_t* full_bar = thr_s + NQ;                                        // [stages]
    uint64_t* empty_bar = full_bar + SE_MAX_STAGES;                         // [stages]
    int* cnt_s = reinterpret_cast<int*>(empty_bar + SE_MAX_STAGES);         // [NQ]

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int64_t num_tiles = (p.n + TR - 1) / TR;

    // ---- set-up ----
    for (int i = tid; i < NQ * p.ld; i += SE_THREADS) {
        int j = i / p.ld, c = i - j * p.ld;
        q_s[i] = (j < p.nq_real && c < p.ldq) ? p.q[(size_t)j * p.ldq + c] : 0.f;
    }
    for (int i = tid; i < NQ * p.cap; i += SE_THREADS) cand[i] = COMP_EMPTY;
    if (tid < NQ) {
        thr_s[tid] = make_comp(FLT_MAX, 0u);
        cnt_s[tid] = 0;
    }
    if (tid == 0) {
        for (int s = 0; s < p.stages; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], SE_CWARPS);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == SE_CWARPS) {
        // ================= producer: one lane streams tiles into the ring =================
        if (lane == 0) {
            int it = 0;
            for (int64_t t = blockIdx.x; t < num_tiles; t += gridDim.x, it++) {
                const int s = it % p.stages;
                const uint32_t ph = (uint32_t)(it / p.stages) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                const int64_t r0 = t * TR;
                const int rows = (int)min((int64_t)TR, p.n - r0);
                const uint32_t bytes = (uint32_t)rows * (uint32_t)row_bytes;
                mbar_expect_tx(&full_bar[s], bytes);
                bulk_g2s(tiles + (size_t)s * tile_bytes,
                         reinterpret_cast<const unsigned char*>(p.xb) + (size_t)r0 * row_bytes, bytes, &full_bar[s]);
            }
        }
        return;
    }

    // ================= consumers =================
    const int nvec = p.ld / VEC;
    int it = 0;
    for (int64_t t = blockIdx.x; t < num_tiles; t += gridDim.x, it++) {
        const int s = it % p.stages;
        const uint32_t ph = (uint32_t)(it / p.stages) & 1u;
        mbar_wait(&full_bar[s], ph);
        const unsigned char* tile = tiles + (size_t)s * tile_bytes + (size_t)(warp * R) * row_bytes;

        float acc[V];
#pragma unroll
        for (int v = 0; v < V; v++) acc[v] = 0.f;
#pragma unroll 3
        for (int c = lane; c < nvec; c += 32) {
            float xv[R][VEC];
#pragma unroll
            for (int i = 0; i < R; i++) RowVec<RowT>::load(tile + (size_t)i * row_bytes + (size_t)c * 16, xv[i]);
#pragma unroll
            for (int j = 0; j < NQ; j++) {
                float qv[VEC];
                const float4* qp = reinterpret_cast<const float4*>(q_s + j * p.ld + c * VEC);
#pragma unroll
                for (int e = 0; e < VEC / 4; e++) {
                    float4 t4 = qp[e];
                    qv[4 * e] = t4.x; qv[4 * e + 1] = t4.y; qv[4 * e + 2] = t4.z; qv[4 * e + 3] = t4.w;
                }
#pragma unroll
                for (int i = 0; i < R; i++) {
#pragma unroll
                    for (int e = 0; e < VEC; e++) {
                        if (IS_L2) {
                            float df = xv[i][e] - qv[e];
                            acc[i * NQ + j] = fmaf(df, df, acc[i * NQ + j]);
                        } else {
                            acc[i * NQ + j] = fmaf(xv[i][e], qv[e], acc[i * NQ + j]);
                        }
                    }
                }
            }
        }
        // this warp is done with the tile's bytes: hand the slot back to the producer
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);

        // transposing butterfly: V values/lane -> 1 value/lane; value v lands on the lanes
        // whose top log2(V) bits equal v
        int cntv = V;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            if (cntv > 1) {
                const bool upper = (lane & off) != 0;
                const int half = cntv >> 1;
#pragma unroll
                for (int i = 0; i < V / 2; i++) {
                    if (i < half) {
                        float keep = upper ? acc[i + half] : acc[i];
                        float send = upper ? acc[i] : acc[i + half];
                        acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                    }
                }
                cntv = half;
            } else {
                acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], off);
            }
        }
        constexpr int LOGV = (V == 1) ? 0 : (V == 2) ? 1 : (V == 4) ? 2 : (V == 8) ? 3 : (V == 16) ? 4 : 5;
        const int v = lane >> (5 - LOGV);
        const bool leader = (lane & ((32 >> LOGV) - 1)) == 0;
        const int ri = v / NQ, qj = v - ri * NQ;
        const int64_t row = t * TR + warp * R + ri;
        if (leader && row < p.n && qj < p.nq_real) {
            const float key = IS_L2 ? acc[0] : -acc[0];
            if (key < FLT_MAX) {  // also rejects NaN, as faiss's strict compare does
                const uint64_t comp = make_comp(key, (uint32_t)row);
                // the selector is only consulted for rows that would enter the list
                if (comp < *reinterpret_cast<volatile uint64_t*>(&thr_s[qj]) &&
                    (p.sel == nullptr || ((p.sel[row >> 3] >> (row & 7)) & 1u))) {
                    int slot = atomicAdd(&cnt_s[qj], 1);
                    if (slot < p.cap) cand[(size_t)qj * p.cap + slot] = comp;
                }
            }
        }

        // ---- prune when a buffer could overflow on the next tile (uniform decision) ----
        consumer_sync();
        int counts[NQ];
        bool any = false;
#pragma unroll
        for (int j = 0; j < NQ; j++) {
            counts[j] = *reinterpret_cast<volatile int*>(&cnt_s[j]);
            any |= (counts[j] + TR > p.cap);
        }
        consumer_sync();  // everyone has read the counts before anyone inserts again
        if (any) {
#pragma unroll
            for (int j = 0; j < NQ; j++) {
                if (counts[j] + TR > p.cap) {
                    uint64_t* cj = cand + (size_t)j * p.cap;
                    consumer_bitonic_sort(cj, p.cap, tid);
                    const int kept = min(counts[j], p.k);
                    for (int i = kept + tid; i < p.cap; i += SE_CTHREADS) cj[i] = COMP_EMPTY;
                    if (tid == 0) {
                        cnt_s[j] = kept;
                        if (kept >= p.k) thr_s[j] = cj[p.k - 1];
                    }
                }
            }
            consumer_sync();
        }
    }

    // ---- drain: sort every query's buffer and publish this CTA's top-k ----
    consumer_sync();
#pragma unroll
    for (int j = 0; j < NQ; j++) {
        if (j < p.nq_real) {
            uint64_t* cj = cand + (size_t)j * p.cap;
            // slots past the count hold COMP_EMPTY: only the filled prefix (rounded up to a power of two that also
            // covers k) needs ordering
            int ns = 32;
            const int filled = *reinterpret_cast<volatile int*>(&cnt_s[j]);
            while (ns < filled || ns < p.k) ns <<= 1;
            consumer_sort(cj, min(ns, p.cap), tid);
            uint64_t* dst = p.part + ((size_t)j * p.nparts + blockIdx.x) * p.k;
            for (int i = tid; i < p.k; i += SE_CTHREADS) dst[i] = cj[i];
        }
    }
    if (p.fuse) {
        __shared__ int last_s;
        __threadfence();  // this CTA's lists, before its ticket
        consumer_sync();
        if (tid == 0) last_s = (atomicAdd(p.ticket, 1) == (int)gridDim.x - 1) ? 1 : 0;
        consumer_sync();
        if (last_s) {
            if (tid == 0) *p.ticket = 0;  // re-armed for the next search
            __threadfence();
            // the ring was last written by bulk copies (async proxy); it is the merge window now
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            uint64_t* buf = reinterpret_cast<uint64_t*>(tiles);
            const int m = p.nparts * p.k;
            for (int j = 0; j < p.nq_real; j++) {
                const uint64_t* src = p.part + (size_t)j * m;
                for (int i = tid; i < p.merge_n; i += SE_CTHREADS) buf[i] = i < m ? __ldcg(src + i) : COMP_EMPTY;
                consumer_sync();
                consumer_sort(buf, p.merge_n, tid);
                for (int i = tid; i < p.k; i += SE_CTHREADS) {
                    const uint64_t c = buf[i];
                    float dv;
                    int64_t iv;
                    if (c == COMP_EMPTY) {  // faiss's padding
                        dv = (p.metric == BFK_METRIC_L2) ? FLT_MAX : -FLT_MAX;
                        iv = -1;
                    } else {
                        const float key = ord2f((uint32_t)(c >> 32));
                        dv = (p.metric == BFK_METRIC_L2) ? key : -key;
                        iv = (int64_t)(c & 0xffffffffu) + p.id_offset;
                    }
                    p.D[(size_t)j * p.out_ld + i] = dv;
                    p.I[(size_t)j * p.out_ld + i] = iv;
                }
                consumer_sync();
            }
            if (p.h_seq != nullptr) {  // (D, I) are pinned host memory and the host polls this word
                __threadfence_system();
                consumer_sync();
                if (tid == 0) {
                    *reinterpret_cast<volatile int*>(p.h_seq) = p.seq;
                    __threadfence_system();
                }
            }
        }
    }
}

// ---- host side ---------------------------------------------------------------
struct ScanExactPlan {
    int R = 4;
    int cap = 0;
    int stages = 0;
    int group = 8;  // queries per launch
    size_t smem = 0;
    int grid = 1;
};

static int plan_scan_exact(const bfk_index* idx, int nq, int k, ScanExactPlan* pl) {
    const bool f32 = idx->storage == BFK_STORE_F32;
    const int ld = f32 ? idx->ld32 : idx->ld16;
    const int row_bytes = ld * (f32 ? 4 : 2);
    int R = 4;
    while (R > 1 && SE_CWARPS * R * row_bytes > 48 * 1024) R >>= 1;
    const int TR = SE_CWARPS * R;
    const int cap = next_pow2(k + 2 * TR);
    // one pass serves up to 8 queries; 3 and 5..7 queries round UP to the next template size (the kernel
    // masks the missing ones) instead of splitting into two corpus passes
    int group = std::min(8, next_pow2(std::max(nq, 1)));
    for (;; group >>= 1) {
        size_t fixed = (size_t)group * ld * 4 + (size_t)group * cap * 8 + group * 8 + 2 * SE_MAX_STAGES * 8 + group * 4 + 128;
        size_t tile = (size_t)TR * row_bytes;
        int stages = (int)std::min<size_t>(SE_MAX_STAGES, (SE_SMEM_BUDGET - std::min<size_t>(fixed, SE_SMEM_BUDGET)) / tile);
        if (stages >= 2 || (stages >= 1 && group == 1)) {
            pl->R = R;
            pl->cap = cap;
            pl->stages = stages;
            pl->group = group;
            pl->smem = fixed + (size_t)stages * tile;
            break;
        }
        if (group == 1) return fail("scan_exact: row too large for shared-memory staging");
    }
    int64_t tiles = (idx->ntotal + TR - 1) / TR;
    pl->grid = (int)std::max<int64_t>(1, std::min<int64_t>(tiles, idx->num_sms));
    return 0;
}

int scan_exact_workspace(const bfk_index* idx, int nq, int k, size_t* bytes, int* nparts_out) {
    ScanExactPlan pl;
    BFK_TRY(plan_scan_exact(idx, std::max(nq, 1), k, &pl));
    *nparts_out = pl.grid;
    *bytes = sizeof(uint64_t) * (size_t)nq * pl.grid * k;
    return 0;
}

template <int NQ, int R, bool IS_L2, typename RowT>
static int launch_one(const ScanExactParams& p, size_t smem, int grid, cudaStream_t st) {
    auto kern = scan_exact_kernel<NQ, R, IS_L2, RowT>;
    static bool attr_set[64] = {false};  // per instantiation and per device
    int dev = 0;
    BFK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        BFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SE_SMEM_BUDGET + 4096));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    kern<<<grid, SE_THREADS, smem, st>>>(p);
    BFK_CUDA(cudaGetLastError());
    return 0;
}

template <int NQ, bool IS_L2, typename RowT>
static int launch_r(int R, const ScanExactParams& p, size_t smem, int grid, cudaStream_t st) {
    switch (R) {
        case 4: return launch_one<NQ, 4, IS_L2, RowT>(p, smem, grid, st);
        case 2: return launch_one<NQ, 2, IS_L2, RowT>(p, smem, grid, st);
        default: return launch_one<NQ, 1, IS_L2, RowT>(p, smem, grid, st);
    }
}

template <bool IS_L2, typename RowT>
static int launch_nq(int NQ, int R, const ScanExactParams& p, size_t smem, int grid, cudaStream_t st) {
    switch (NQ) {
        case 8: return launch_r<8, IS_L2, RowT>(R, p, smem, grid, st);
        case 4: return launch_r<4, IS_L2, RowT>(R, p, smem, grid, st);
        case 2: return launch_r<2, IS_L2, RowT>(R, p, smem, grid, st);
        default: return launch_r<1, IS_L2, RowT>(R, p, smem, grid, st);
    }
}

// nq <= 8 queries; part is [nq][nparts][k]
int launch_scan_exact(const bfk_index* idx, const float* q_dev, int ldq, int nq, int k, uint64_t* part, int nparts,
                      cudaStream_t st, const unsigned char* sel, const ScanFuse* fuse, bool* fused) {
    if (fused) *fused = false;
    ScanExactPlan pl;
    BFK_TRY(plan_scan_exact(idx, nq, k, &pl));
    if (pl.grid != nparts) {
        // the caller sized `part` for a (possibly) different group size; grids only depend on
        // ntotal and the row size, so this is an internal inconsistency
        return fail("internal: scan_exact grid/nparts mismatch");
    }
    const bool f32 = idx->storage == BFK_STORE_F32;
    for (int q0 = 0; q0 < nq;) {
        int g = std::min(nq - q0, pl.group);
        int NQ = next_pow2(g);
        ScanExactPlan sub;
        BFK_TRY(plan_scan_exact(idx, NQ, k, &sub));
        if (sub.group < NQ) {  // shared memory cannot hold NQ candidate buffers: shrink the group
            NQ = sub.group;
            g = std::min(g, NQ);
        }
        ScanExactParams p;
        p.xb = f32 ? (const void*)idx->x32 : (const void*)idx->x16;
        p.n = idx->ntotal;
        p.ld = f32 ? idx->ld32 : idx->ld16;
        p.q = q_dev + (size_t)q0 * ldq;
        p.ldq = ldq;
        p.nq_real = g;
        p.k = k;
        p.cap = sub.cap;
        p.stages = sub.stages;
        p.part = part + (size_t)q0 * nparts * k;
        p.nparts = nparts;
        p.sel = sel;
        p.fuse = 0;
        p.ticket = nullptr;
        p.h_seq = nullptr;
        if (fuse && q0 == 0 && g == nq) {
            // one launch covers the search: its last CTA can merge, if the lists are few enough for one CTA to sort
            // quickly (beyond that K4's one-CTA-per-query launch is faster) and the window fits the ring
            int n = 32;
            while (n < nparts * k || n < k) n <<= 1;
            const size_t ring = (size_t)sub.stages * (size_t)(SE_CWARPS * sub.R) * (size_t)p.ld * (f32 ? 4 : 2);
            if ((int64_t)nq * n <= 2048 && (size_t)n * 8 <= ring) {
                p.fuse = 1;
                p.metric = idx->metric;
                p.merge_n = n;
                p.id_offset = fuse->id_offset;
                p.D = fuse->D;
                p.I = fuse->I;
                p.out_ld = fuse->out_ld;
                p.ticket = fuse->ticket;
                p.h_seq = fuse->h_seq;
                p.seq = fuse->seq;
                if (fused) *fused = true;
            }
        }
        const bool l2 = idx->metric == BFK_METRIC_L2;
        int rc;
        if (f32)
            rc = l2 ? launch_nq<true, float>(NQ, sub.R, p, sub.smem, sub.grid, st)
                    : launch_nq<false, float>(NQ, sub.R, p, sub.smem, sub.grid, st);
        else
            rc = l2 ? launch_nq<true, __nv_bfloat16>(NQ, sub.R, p, sub.smem, sub.grid, st)
                    : launch_nq<false, __nv_bfloat16>(NQ, sub.R, p, sub.smem, sub.grid, st);
        BFK_TRY(rc);
        q0 += g;
    }
    return 0;
}

}  // namespace bfk

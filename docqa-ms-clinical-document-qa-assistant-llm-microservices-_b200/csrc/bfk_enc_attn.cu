// E1 `enc_embed_ln_kernel`, E3 attention: the non-GEMM parts of the sentence encoder
// (transformers BertEmbeddings and BertSelfAttention as run by
// SentenceTransformer('all-MiniLM-L6-v2') for semantic-indexer/indexer.py:21,37 and llm-qa/main.py:25).
//
// Four attention kernels live here, in the order they were written; the last one ships, the others stay selectable
// (BFK_ENC_ATTN=row|blk|ws) and tested, as the measured steps between them:
//   enc_attention_kernel      whole 128 x T score tile in TMEM                       465 us per layer (512 chunks x 256 tokens)
//   enc_attention_blk_kernel  64-key blocks, four CTAs per SM, one softmax shift      344 -> 287 us
//   enc_attention_ws_kernel   warp-specialised roles over mbarriers, TMA staging     217 us
//   enc_attention_pws_kernel  the same in persistent CTAs with the next item prefetched  198 us
//
// enc_attention_kernel: one CTA per (128-query tile, head, chunk), 8 warps, two threads per query row (row == TMEM lane; each
// owns half of the key columns and exchanges its row max / row sum with its partner through shared memory):
//   1. Q [128 x 32], K [T x 32] and V^T [32 x T] of this head are written into shared memory in the
//      128-byte-swizzled K-major layout tcgen05 reads (the rows are 64 bytes, so the upper half of each
//      swizzle row is simply unused);
//   2. S = Q K^T : tcgen05.mma M=128, N=T, two K=16 steps, fp32 accumulator in TMEM (T <= 256 columns);
//   3. softmax over the row, in the thread that owns it: pass 1 max of (S/sqrt(32) + mask bias), pass 2
//      exp2 and sum; P is rounded to bf16 and written to shared memory as the A operand of step 4 (the
//      region Q and K occupied);
//   4. O = P V : T/64 K-blocks x 4 tcgen05.mma (M=128, N=32, K=16) into the first 32 TMEM columns;
//   5. O / rowsum -> bf16 context rows.
// 256 TMEM columns and ~82 KB of shared memory per CTA: two CTAs per SM overlap each other's phases.
#include <cstdlib>

#include "bfk_common.cuh"
#include "bfk_ptx.cuh"
#include "bfk_tmap.cuh"

namespace bfk {

constexpr int EA_THREADS = 256;      // two threads per query row: (row = TMEM lane, half of the keys)
constexpr int EA_DH = 32;        // head dimension (all-MiniLM-L6-v2: 384 / 12)
constexpr int EA_MAXT = 256;
constexpr int EA_Q_OFF = 0;                   // 16 KiB
constexpr int EA_K_OFF = 16 * 1024;           // 32 KiB
constexpr int EA_P_OFF = 0;                   // 64 KiB, aliases Q and K once S is in TMEM
constexpr int EA_VT_OFF = 64 * 1024;          // 16 KiB
constexpr int EA_MB_OFF = 80 * 1024;          // [256] fp32 mask bias
constexpr int EA_BAR_OFF = EA_MB_OFF + EA_MAXT * 4;
constexpr int EA_RED_OFF = EA_BAR_OFF + 64;   // [2 halves][128 rows] row max, then row sums
constexpr int EA_SMEM = EA_RED_OFF + 2 * 2 * 128 * 4 + 1024 /*alignment*/;

// 2^x on the SFU without exp2f's denormal-range fix-up (a probability below 2^-126 is zero in bf16 anyway)
__device__ __forceinline__ float ex2_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ uint32_t pack_bf16x2_rn(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}

__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// byte offset of 16-byte chunk `c` (0..7) of row `r` inside a [rows x 128 B] tile, 128-byte swizzle
__device__ __forceinline__ uint32_t sw128_off(int r, int c) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}

struct EncAttnParams {
    const __nv_bfloat16* qkv;   // [B*T, 3*H]: Q | K | V, head h at columns h*32
    const float* mask_bias;     // [B*T]: 0 for real tokens, -1e30 for padding (already in the log2 domain)
    __nv_bfloat16* ctx;         // [B*T, H]
    int T, H;
    float scale_log2;           // log2(e) / sqrt(head_dim)
    float shift_limit;          // the bound ||q|| max||k|| replaces the row maximum while it is <= this (60; 0 = never)
    unsigned long long* trace;  // debugging (BFK_ENC_ATTN_TRACE): [128] clock64 stamps of one CTA, or null
};

__global__ void __launch_bounds__(EA_THREADS) enc_attention_kernel(const EncAttnParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    unsigned char* smem = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    float* mb = reinterpret_cast<float*>(smem + EA_MB_OFF);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + EA_BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
    float* red_max = reinterpret_cast<float*>(smem + EA_RED_OFF);  // [2][128]
    float* red_sum = red_max + 256;                                // [2][128]

    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = tid & 127, half = tid >> 7;
    const int mt = blockIdx.x, h = blockIdx.y;
    const int64_t b = blockIdx.z;
    const int T = p.T, ld = 3 * p.H;
    const int64_t row0 = b * T;

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc<256>(tmem_slot);

    // ---- 1. stage Q, K, V^T and the mask bias ----
    {
        const int q = mt * 128 + row;
        uint4 v[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
        if (q < T) {
            const uint4* src = reinterpret_cast<const uint4*>(p.qkv + (size_t)(row0 + q) * ld + h * EA_DH) + 2 * half;
            v[0] = __ldg(src);
            v[1] = __ldg(src + 1);
        }
        *reinterpret_cast<uint4*>(smem + EA_Q_OFF + sw128_off(row, 2 * half)) = v[0];
        *reinterpret_cast<uint4*>(smem + EA_Q_OFF + sw128_off(row, 2 * half + 1)) = v[1];
    }
    for (int t = tid; t < T; t += EA_THREADS) {
        const uint4* ksrc = reinterpret_cast<const uint4*>(p.qkv + (size_t)(row0 + t) * ld + p.H + h * EA_DH);
        const uint4* vsrc = reinterpret_cast<const uint4*>(p.qkv + (size_t)(row0 + t) * ld + 2 * p.H + h * EA_DH);
        uint4 kv[4], vv[4];
#pragma unroll
        for (int c = 0; c < 4; c++) {
            kv[c] = __ldg(ksrc + c);
            vv[c] = __ldg(vsrc + c);
        }
#pragma unroll
        for (int c = 0; c < 4; c++) *reinterpret_cast<uint4*>(smem + EA_K_OFF + sw128_off(t, c)) = kv[c];
        // V^T: element (dim, token) -> K-block token/64, row dim, column token%64
        const unsigned short* ve = reinterpret_cast<const unsigned short*>(vv);
        const int kb = t >> 6, col = t & 63;
        unsigned char* vt = smem + EA_VT_OFF + kb * 4096;
#pragma unroll
        for (int d = 0; d < EA_DH; d++)
            *reinterpret_cast<unsigned short*>(vt + sw128_off(d, col >> 3) + (col & 7) * 2) = ve[d];
        mb[t] = p.mask_bias[row0 + t];
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

    // ---- 2. S = Q K^T ----
    if (warp == 0) {
        if (elect_one()) {
            const uint32_t idesc = make_idesc_bf16(128, T);
            const uint64_t da0 = make_smem_desc_sw128(smem_u32(smem + EA_Q_OFF));
            const uint64_t db0 = make_smem_desc_sw128(smem_u32(smem + EA_K_OFF));
#pragma unroll
            for (int kk = 0; kk < EA_DH / 16; kk++) umma_bf16(tmem_base, da0 + 2 * kk, db0 + 2 * kk, idesc, kk != 0 ? 1u : 0u);
            umma_commit(&bars[0]);
        }
        __syncwarp();
    }
    mbar_wait(&bars[0], 0);
    tc_fence_after();

    // ---- 3. softmax of row `row`: this thread owns the key columns [c_lo, c_lo + T/2) ----
    const uint32_t lane_addr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const int c_lo = half * (T >> 1), nch = T >> 6;  // 32-column chunks per thread
    uint32_t r[2][32];
    float mx = -INFINITY;
    tmem_ld_32x32(lane_addr + c_lo, r[0]);
#pragma unroll 1
    for (int ch = 0; ch < nch; ch += 2) {  // two register sets: the next chunk's load is in flight
        tmem_ld_wait();
        if (ch + 1 < nch) tmem_ld_32x32(lane_addr + c_lo + (ch + 1) * 32, r[1]);
#pragma unroll
        for (int j = 0; j < 32; j++) mx = fmaxf(mx, fmaf(__uint_as_float(r[0][j]), p.scale_log2, mb[c_lo + ch * 32 + j]));
        if (ch + 1 < nch) {
            tmem_ld_wait();
            if (ch + 2 < nch) tmem_ld_32x32(lane_addr + c_lo + (ch + 2) * 32, r[0]);
#pragma unroll
            for (int j = 0; j < 32; j++)
                mx = fmaxf(mx, fmaf(__uint_as_float(r[1][j]), p.scale_log2, mb[c_lo + (ch + 1) * 32 + j]));
        }
    }
    red_max[half * 128 + row] = mx;
    tmem_ld_32x32(lane_addr + c_lo, r[0]);  // pass 2's first chunk loads while the halves exchange their maxima
    __syncthreads();
    mx = fmaxf(mx, red_max[(half ^ 1) * 128 + row]);
    // P may overwrite Q and K: every row's S is complete (bars[0]) and nobody reads Q / K from shared memory
    float sum = 0.f;
    auto exp32 = [&](const uint32_t (&rv)[32], int c) {  // c = first key column of the chunk
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const float p0 = ex2_fast(fmaf(__uint_as_float(rv[2 * j]), p.scale_log2, mb[c + 2 * j]) - mx);
            const float p1 = ex2_fast(fmaf(__uint_as_float(rv[2 * j + 1]), p.scale_log2, mb[c + 2 * j + 1]) - mx);
            const __nv_bfloat162 pb = __floats2bfloat162_rn(p0, p1);
            sum += __bfloat162float(pb.x) + __bfloat162float(pb.y);  // normalise by what is really multiplied
            pk[j] = *reinterpret_cast<const uint32_t*>(&pb);
        }
        unsigned char* pblk = smem + EA_P_OFF + (c >> 6) * 16384;
#pragma unroll
        for (int i = 0; i < 4; i++)
            *reinterpret_cast<uint4*>(pblk + sw128_off(row, ((c & 63) >> 3) + i)) =
                make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
    };
#pragma unroll 1
    for (int ch = 0; ch < nch; ch += 2) {
        tmem_ld_wait();
        if (ch + 1 < nch) tmem_ld_32x32(lane_addr + c_lo + (ch + 1) * 32, r[1]);
        exp32(r[0], c_lo + ch * 32);
        if (ch + 1 < nch) {
            tmem_ld_wait();
            if (ch + 2 < nch) tmem_ld_32x32(lane_addr + c_lo + (ch + 2) * 32, r[0]);
            exp32(r[1], c_lo + (ch + 1) * 32);
        }
    }
    red_sum[half * 128 + row] = sum;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    sum += red_sum[(half ^ 1) * 128 + row];

    // ---- 4. O = P V ----
    if (warp == 0) {
        if (elect_one()) {
            const uint32_t idesc = make_idesc_bf16(128, EA_DH);
            for (int kb = 0; kb < (T >> 6); kb++) {
                const uint64_t da0 = make_smem_desc_sw128(smem_u32(smem + EA_P_OFF + kb * 16384));
                const uint64_t db0 = make_smem_desc_sw128(smem_u32(smem + EA_VT_OFF + kb * 4096));
#pragma unroll
                for (int kk = 0; kk < 4; kk++) umma_bf16(tmem_base, da0 + 2 * kk, db0 + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
            }
            umma_commit(&bars[1]);
        }
        __syncwarp();
    }
    mbar_wait(&bars[1], 0);
    tc_fence_after();

    // ---- 5. normalise and store: this thread writes dims [16 * half, +16) of its row ----
    uint32_t ro[16];
    tmem_ld_32x16(lane_addr + half * 16, ro);
    tmem_ld_wait();
    const int q = mt * 128 + row;
    if (q < T) {
        const float inv = 1.0f / sum;
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const __nv_bfloat162 ob = __floats2bfloat162_rn(__uint_as_float(ro[2 * j]) * inv, __uint_as_float(ro[2 * j + 1]) * inv);
            o[j] = *reinterpret_cast<const uint32_t*>(&ob);
        }
        uint4* dst = reinterpret_cast<uint4*>(p.ctx + (size_t)(row0 + q) * p.H + h * EA_DH + half * 16);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

// ---- E3b: the same attention in 64-key blocks, four CTAs per SM -------------------------------------------
// enc_attention_kernel above keeps a whole 128 x T score tile in TMEM (256 columns): two CTAs per SM, each one
// serial latency chain (ncu: 24 % of the warps active, 465 us per layer).  This variant needs 128 TMEM columns
// and 55 KB of shared memory per CTA, so four chains overlap on an SM.  One CTA per (head, chunk): K is staged
// once for the chunk's one or two 128-query tiles; thread == query row == TMEM lane.
//   pass 1  (only when the cheap bound on the row maximum is unusable) S = Q K^T in 128-key pieces for the maximum;
//   pass 2  per 64-key block j: S_j into columns 0..63, p = exp2(s - max) in the row's thread, P_j (bf16) and
//           V_j^T staged in shared memory, O (+)= P_j V_j into columns 64..111.  S_{j+1} is issued together with
//           P_j V_j, so it is complete by the time the threads come back for it.
// The row sums come out of the same MMA: V^T carries a 33rd row of ones (N = 48), so column 96 of O is
// sum_j p_j -- of the bf16-rounded p that are really multiplied.  A block whose keys are all masked is skipped
// (its p are zero); a block without masked keys needs no per-key test.
// Q and K use the 64-byte swizzle (their rows are 64 bytes: head dimension 32), P and V^T the 128-byte one.
// (Two threads per row -- 256-thread CTAs, eight warps per scheduler -- was measured slower: 372 us against 287 us
// per layer of 512 chunks x 256 tokens.)
constexpr int EB_THREADS = 128;
constexpr int EB_ON = 48;                   // N of the P V product: 32 dims, the ones row, padding to a multiple of 16
constexpr int EB_Q_OFF = 0;                 // [2 tiles][128 x 64 B]  16 KiB
constexpr int EB_K_OFF = 16 * 1024;         // [T x 64 B]          <= 16 KiB
constexpr int EB_V_OFF = 32 * 1024;         // [48 rows x 128 B]       6 KiB (one 64-token block; rows 32..47 constant)
constexpr int EB_P_OFF = 38 * 1024;         // [128 x 128 B]          16 KiB (one 64-key block)
constexpr int EB_BAR_OFF = 54 * 1024;
constexpr int EB_SMEM = EB_BAR_OFF + 128;

// byte offset of 16-byte chunk `c` (0..3) of row `r` inside a [rows x 64 B] tile, 64-byte swizzle
__device__ __forceinline__ uint32_t sw64_off(int r, int c) {
    return (uint32_t)((r >> 3) * 512 + (r & 7) * 64 + ((c ^ ((r >> 1) & 3)) << 4));
}
// K-major operand descriptor, 64-byte swizzle: 8-row groups 512 bytes apart
__device__ __forceinline__ uint64_t make_smem_desc_sw64(uint32_t smem_addr_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr_bytes & 0x3ffffu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(512u >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)4 << 61;
    return d;
}

#define EB_MARK(ev) do { if (trace) trace[(ev)] = (unsigned long long)clock64(); } while (0)

__global__ void __launch_bounds__(EB_THREADS, 4) enc_attention_blk_kernel(const EncAttnParams p) {
    // four CTAs per SM leave no room for alignment slack: the kernel has no static shared memory, so the dynamic
    // window starts on a 1 KiB boundary (the swizzled operand tiles need it); anything else is refused
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = smem_raw;
    if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + EB_BAR_OFF);  // [0] scores ready, [1] O block accumulated
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
    float* kmax_s = reinterpret_cast<float*>(tmem_slot + 2);  // [4] one per warp
    uint32_t* keep_s = tmem_slot + 8;                         // [8] bit t%32 of word t/32: key t is not masked

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h = blockIdx.x;
    const int64_t b = blockIdx.y;
    const int T = p.T, ld = 3 * p.H, NB = T >> 6, QT = (T + 127) >> 7;
    const int64_t row0 = b * T;
    // BFK_ENC_ATTN_TRACE: thread 0 of one CTA in the middle of the grid records clock64 at every phase boundary
    unsigned long long* trace = (p.trace != nullptr && tid == 0 && blockIdx.x == 5 && blockIdx.y == gridDim.y / 2) ? p.trace : nullptr;
    EB_MARK(0);

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc<128>(tmem_slot);
    EB_MARK(1);

    // ---- stage Q (this thread's row of each 128-query tile), K (all T rows) and the mask: every load is issued
    //      before the first one is used (one round trip to L2 / HBM instead of three).  On the way, ||q||^2 of the
    //      rows and the largest ||k||^2 of the head: every score obeys s <= ||q|| max||k|| (Cauchy-Schwarz) ----
    auto sumsq = [](const uint4 (&w)[4]) {
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint32_t e[4] = {w[c].x, w[c].y, w[c].z, w[c].w};
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float lo = __uint_as_float(e[i] << 16), hi = __uint_as_float(e[i] & 0xffff0000u);
                s = fmaf(lo, lo, s);
                s = fmaf(hi, hi, s);
            }
        }
        return s;
    };
    float qn2[2] = {0.f, 0.f}, kn2 = 0.f;
    {
        uint4 qv[2][4], kv[2][4];
        float bias[2] = {-1e30f, -1e30f};
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const int t = tid + i * EB_THREADS;  // T is a multiple of 64: t < T is warp-uniform
#pragma unroll
            for (int c = 0; c < 4; c++) qv[i][c] = kv[i][c] = make_uint4(0, 0, 0, 0);
            if (t < T) {
                const uint4* qsrc = reinterpret_cast<const uint4*>(p.qkv + (size_t)(row0 + t) * ld + h * EA_DH);
                const uint4* ksrc = reinterpret_cast<const uint4*>(p.qkv + (size_t)(row0 + t) * ld + p.H + h * EA_DH);
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    qv[i][c] = __ldg(qsrc + c);
                    kv[i][c] = __ldg(ksrc + c);
                }
                bias[i] = p.mask_bias[row0 + t];
            }
        }
        EB_MARK(2);
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const int t = tid + i * EB_THREADS;
            if (i < QT) {  // rows past T inside the last tile are zero
                qn2[i] = sumsq(qv[i]);
#pragma unroll
                for (int c = 0; c < 4; c++) *reinterpret_cast<uint4*>(smem + EB_Q_OFF + i * 8192 + sw64_off(tid, c)) = qv[i][c];
            }
            if (t < T) {
                kn2 = fmaxf(kn2, sumsq(kv[i]));
#pragma unroll
                for (int c = 0; c < 4; c++) *reinterpret_cast<uint4*>(smem + EB_K_OFF + sw64_off(t, c)) = kv[i][c];
                const uint32_t keep = __ballot_sync(0xffffffffu, bias[i] == 0.f);
                if (lane == 0) keep_s[t >> 5] = keep;
            }
        }
    }
    EB_MARK(3);
    // rows 32..47 of V^T never change: row 32 is all ones (the row sums), the rest zero; a row is 128 contiguous
    // bytes whatever the swizzle does inside it
    {
        const uint32_t w = tid < 8 ? 0x3f803f80u : 0u;
        *reinterpret_cast<uint4*>(smem + EB_V_OFF + 32 * 128 + tid * 16) = make_uint4(w, w, w, w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kn2 = fmaxf(kn2, __shfl_xor_sync(0xffffffffu, kn2, o));
    if (lane == 0) kmax_s[warp] = kn2;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    EB_MARK(4);
    kn2 = fmaxf(fmaxf(kmax_s[0], kmax_s[1]), fmaxf(kmax_s[2], kmax_s[3]));
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
    const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
    uint32_t ph_s = 0, ph_o = 0;  // phases of bars[0] / bars[1]
    uint32_t r[2][32];

    // which 64-key blocks take part, and which of them have no masked key (CTA-uniform)
    uint32_t act = 0, full = 0;
    for (int j = 0; j < NB; j++) {
        const uint32_t w0 = keep_s[2 * j], w1 = keep_s[2 * j + 1];
        if (w0 | w1) act |= 1u << j;
        if ((w0 & w1) == 0xffffffffu) full |= 1u << j;
    }
    // A chunk whose every key is masked: all its scores carry the same -1e30 bias, which swallows them in fp32, and
    // the softmax is uniform over all T keys.  Scale 0 and shift 0 over all blocks give exactly that.
    const bool no_key = act == 0;
    if (no_key) act = full = (1u << NB) - 1u;
    const float scale = no_key ? 0.f : p.scale_log2;

    // ---- the softmax shift ----
    // Any mx >= the row's largest score gives the same softmax; it need not be the maximum itself.  The bound
    // ||q|| max||k|| (in the exp2 domain) costs nothing, and p = exp2(s - mx) keeps its relative precision however
    // far mx lies above the true maximum -- fp32 and bf16 share their exponent range -- as long as nothing
    // underflows: |s| <= mx, so with the bound at most 60 every p is >= 2^-120, a normal number in both formats,
    // and the first pass over the keys, a quarter of this kernel, is skipped.  CTAs with a larger bound take the
    // exact maximum as before.
    const float mhat0 = sqrtf(qn2[0] * kn2) * scale * 1.0001f + 1e-3f;
    const float mhat1 = sqrtf(qn2[1] * kn2) * scale * 1.0001f + 1e-3f;
    const bool exact_max = __syncthreads_or(fmaxf(mhat0, mhat1) > p.shift_limit ? 1 : 0) != 0 && !no_key;

    const uint32_t idesc_s = make_idesc_bf16(128, 64), idesc_o = make_idesc_bf16(128, EB_ON);
    const int tt = tid >> 1, hf = tid & 1;  // V staging: token of the block, 16-dim half
    unsigned char* Pb = smem + EB_P_OFF;
    unsigned char* Vb = smem + EB_V_OFF;
    auto issue_scores = [&](int mt, int j) {  // warp 0, converged: S_j = Q_mt K_j^T into columns 0..63
        if (elect_one()) {
            const uint64_t dq = make_smem_desc_sw64(smem_u32(smem + EB_Q_OFF + mt * 8192));
            const uint64_t dk = make_smem_desc_sw64(smem_u32(smem + EB_K_OFF + j * 64 * 64));
#pragma unroll
            for (int kk = 0; kk < EA_DH / 16; kk++) umma_bf16(tmem_base, dq + 2 * kk, dk + 2 * kk, idesc_s, kk != 0 ? 1u : 0u);
            umma_commit(&bars[0]);
        }
        __syncwarp();
    };
    const int j_first = __ffs(act) - 1;
    if (!exact_max && warp == 0) issue_scores(0, j_first);
    EB_MARK(5);

    for (int mt = 0; mt < QT; mt++) {
        float mx = exact_max ? -INFINITY : (mt == 0 ? mhat0 : mhat1);
        // ---- pass 1 (only when the bound is not usable): row maximum over the kept keys, 128 keys at a time ----
        for (int k0 = 0; exact_max && k0 < T; k0 += 128) {
            const int n1 = min(128, T - k0);
            if (warp == 0) {
                if (elect_one()) {
                    const uint32_t idesc = make_idesc_bf16(128, n1);
                    const uint64_t dq = make_smem_desc_sw64(smem_u32(smem + EB_Q_OFF + mt * 8192));
                    const uint64_t dk = make_smem_desc_sw64(smem_u32(smem + EB_K_OFF + k0 * 64));
#pragma unroll
                    for (int kk = 0; kk < EA_DH / 16; kk++) umma_bf16(tmem_base, dq + 2 * kk, dk + 2 * kk, idesc, kk != 0 ? 1u : 0u);
                    umma_commit(&bars[0]);
                }
                __syncwarp();
            }
            mbar_wait(&bars[0], ph_s);
            ph_s ^= 1u;
            tc_fence_after();
            for (int c = 0; c < n1; c += 32) {
                tmem_ld_32x32(lane_addr + c, r[0]);
                tmem_ld_wait();
                const uint32_t w = keep_s[(k0 + c) >> 5];
#pragma unroll
                for (int j = 0; j < 32; j++)
                    if ((w >> j) & 1u) mx = fmaxf(mx, __uint_as_float(r[0][j]) * scale);
            }
            tc_fence_before();
            __syncthreads();  // every row has read this piece before the next MMA overwrites the columns
            tc_fence_after();
        }
        if (exact_max && warp == 0) issue_scores(mt, j_first);

        // ---- pass 2: per 64-key block, p = exp2(s - max), O += P_j V_j ----
        const float neg_mx = -mx;
        bool first = true;
        int ev = 8 + mt * 32;
        for (int j = 0; j < NB; j++) {
            if (!((act >> j) & 1u)) continue;
            EB_MARK(ev + 0);
            // this block's V rows travel to registers while the MMA runs
            const uint4* vsrc =
                reinterpret_cast<const uint4*>(p.qkv + (size_t)(row0 + j * 64 + tt) * ld + 2 * p.H + h * EA_DH + hf * 16);
            uint4 vv[2];
            vv[0] = __ldg(vsrc);
            vv[1] = __ldg(vsrc + 1);
            mbar_wait(&bars[0], ph_s);
            ph_s ^= 1u;
            tc_fence_after();
            EB_MARK(ev + 1);
            tmem_ld_32x32(lane_addr, r[0]);
            tmem_ld_32x32(lane_addr + 32, r[1]);
            tmem_ld_wait();
            EB_MARK(ev + 2);
            // P and V^T of the previous block are free once its P V product has completed (a new tile starts after
            // the previous tile's last product has been waited for)
            if (!first) {
                mbar_wait(&bars[1], ph_o);
                ph_o ^= 1u;
            }
            EB_MARK(ev + 3);
            const bool unmasked = (full >> j) & 1u;
#pragma unroll
            for (int hh = 0; hh < 2; hh++) {
                uint32_t pk[16];
                if (unmasked) {
#pragma unroll
                    for (int jj = 0; jj < 16; jj++) {
                        const float p0 = ex2_fast(fmaf(__uint_as_float(r[hh][2 * jj]), scale, neg_mx));
                        const float p1 = ex2_fast(fmaf(__uint_as_float(r[hh][2 * jj + 1]), scale, neg_mx));
                        pk[jj] = pack_bf16x2_rn(p0, p1);
                    }
                } else {
                    const uint32_t w = keep_s[2 * j + hh];
#pragma unroll
                    for (int jj = 0; jj < 16; jj++) {
                        const float p0 = (w >> (2 * jj)) & 1u ? ex2_fast(fmaf(__uint_as_float(r[hh][2 * jj]), scale, neg_mx)) : 0.f;
                        const float p1 = (w >> (2 * jj + 1)) & 1u ? ex2_fast(fmaf(__uint_as_float(r[hh][2 * jj + 1]), scale, neg_mx)) : 0.f;
                        pk[jj] = pack_bf16x2_rn(p0, p1);
                    }
                }
#pragma unroll
                for (int i = 0; i < 4; i++)
                    *reinterpret_cast<uint4*>(Pb + sw128_off(tid, hh * 4 + i)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
            }
            EB_MARK(ev + 4);
            {
                const unsigned short* ve = reinterpret_cast<const unsigned short*>(vv);
#pragma unroll
                for (int e = 0; e < 16; e++)
                    *reinterpret_cast<unsigned short*>(Vb + sw128_off(hf * 16 + e, tt >> 3) + (tt & 7) * 2) = ve[e];
            }
            EB_MARK(ev + 5);
            fence_proxy_async_smem();
            tc_fence_before();
            __syncthreads();  // every row has its scores in registers and P_j, V_j^T in shared memory
            tc_fence_after();
            EB_MARK(ev + 6);
            if (warp == 0) {
                // the next scores first (the threads wait for them next): the next block of this tile, or the first
                // block of the next tile when its shift needs no pass 1
                const uint32_t rest = act >> (j + 1);
                if (rest) issue_scores(mt, j + __ffs(rest));
                else if (!exact_max && mt + 1 < QT) issue_scores(mt + 1, j_first);
                if (elect_one()) {
                    const uint64_t dp = make_smem_desc_sw128(smem_u32(Pb));
                    const uint64_t dv = make_smem_desc_sw128(smem_u32(Vb));
#pragma unroll
                    for (int kk = 0; kk < 4; kk++)
                        umma_bf16(tmem_base + 64, dp + 2 * kk, dv + 2 * kk, idesc_o, (!first || kk != 0) ? 1u : 0u);
                    umma_commit(&bars[1]);
                }
                __syncwarp();
            }
            first = false;
            ev += 7;
        }
        EB_MARK(ev);
        mbar_wait(&bars[1], ph_o);
        ph_o ^= 1u;
        tc_fence_after();
        EB_MARK(ev + 1);

        // ---- normalise and store: columns 64..95 = O, column 96 = the row sum ----
        uint32_t rs[16];
        tmem_ld_32x32(lane_addr + 64, r[0]);
        tmem_ld_32x16(lane_addr + 96, rs);
        tmem_ld_wait();
        const int q = mt * 128 + tid;
        if (q < T) {
            const float inv = 1.0f / __uint_as_float(rs[0]);
            uint32_t o[16];
#pragma unroll
            for (int jj = 0; jj < 16; jj++)
                o[jj] = pack_bf16x2_rn(__uint_as_float(r[0][2 * jj]) * inv, __uint_as_float(r[0][2 * jj + 1]) * inv);
            uint4* dst = reinterpret_cast<uint4*>(p.ctx + (size_t)(row0 + q) * p.H + h * EA_DH);
#pragma unroll
            for (int i = 0; i < 4; i++) dst[i] = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
        }
        tc_fence_before();
        __syncthreads();  // O has been read by every row before the next tile's first product overwrites it
        tc_fence_after();
        EB_MARK(ev + 2);
    }
    if (warp == 0) tmem_dealloc<128>(tmem_base);
    EB_MARK(7);
}


// ---- E3c: warp-specialised attention ---------------------------------------------------------------------
// The timeline of one enc_attention_blk_kernel CTA (BFK_ENC_ATTN_TRACE) showed 4.4 k cycles per 64-key block of
// which the 64 exponentials are 1.1-2.5 k: the rest is the row threads themselves waiting for the scores, issuing
// the MMAs (~1 k cycles in warp 0), staging V^T with 2-byte stores and meeting at a CTA-wide barrier.  Here the
// row threads only do: scores from TMEM -> exp2 -> P into shared memory.  One CTA per (head, chunk), 10 warps:
//   warps 0-3 / 4-7  softmax group A / B: the chunk's first / second 128-query tile, thread == row == TMEM lane
//   warp 8           issues every MMA (scores of the next block as soon as a group has read the current ones)
//   warp 9           stages V: 64-token blocks in their natural [token][32 dims] layout, double-buffered -- the
//                    P V product reads V as an MN-major operand, so nothing is transposed
// and mbarriers between them instead of CTA-wide barriers:
//   s_full[g]   scores of the group's next block are in TMEM            (tcgen05.commit)
//   s_free[g]   the group's 128 threads hold them in registers          -> the next scores may be issued
//   p_full[g]   the group's P block is in shared memory                 -> P V and the row sums may be issued
//   pv_done[g]  that product has completed                              (tcgen05.commit) -> P may be overwritten
//   v_full[s] / v_free[s]  V slot s staged / consumed by the products of both groups
// TMEM: 256 columns (per group: scores 0..63, O 64..95, row sums 96..111), 75 KB of shared memory: two CTAs,
// i.e. four query tiles, per SM as before.  The row sums are a second small product P x ones (N = 16).
constexpr int EW_THREADS = 320;
constexpr int EW_Q_OFF = 0;                 // [2 tiles][128 x 64 B]       16 KiB, 64-byte swizzle
constexpr int EW_K_OFF = 16 * 1024;         // [T x 64 B]               <= 16 KiB, 64-byte swizzle
constexpr int EW_V_OFF = 32 * 1024;         // [2 slots][64 tokens x 64 B]  8 KiB, 64-byte swizzle
constexpr int EW_P_OFF = 40 * 1024;         // [2 groups][128 x 128 B]     32 KiB, 128-byte swizzle
constexpr int EW_ONE_OFF = 72 * 1024;       // [16 rows x 128 B]            2 KiB: row 0 = ones
constexpr int EW_BAR_OFF = 74 * 1024;
constexpr int EW_SMEM = EW_BAR_OFF + 256 + 1024 /*alignment*/;
enum { EW_S_FULL = 0, EW_S_FREE = 2, EW_P_FULL = 4, EW_PV_DONE = 6, EW_V_FULL = 8, EW_V_FREE = 10, EW_P1 = 12, EW_KQ = 13, EW_NBAR = 14 };

__global__ void __launch_bounds__(EW_THREADS, 2)
enc_attention_ws_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmCTX, const EncAttnParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    unsigned char* smem = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + EW_BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + EW_NBAR + 1);
    float* kmax_s = reinterpret_cast<float*>(tmem_slot + 2);  // [8] one per warp
    uint32_t* keep_s = tmem_slot + 12;                        // [8] bit t%32 of word t/32: key t is not masked

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h = blockIdx.x;
    const int64_t b = blockIdx.y;
    const int T = p.T, NB = T >> 6, QT = (T + 127) >> 7;
    const int64_t row0 = b * T;

    if (tid == 0) {
        for (int g = 0; g < 2; g++) {
            mbar_init(&bars[EW_S_FULL + g], 1);
            mbar_init(&bars[EW_S_FREE + g], 128);
            mbar_init(&bars[EW_P_FULL + g], 128);
            mbar_init(&bars[EW_PV_DONE + g], 1);
            mbar_init(&bars[EW_V_FULL + g], 1);
            mbar_init(&bars[EW_V_FREE + g], 1);
        }
        mbar_init(&bars[EW_P1], 1);
        mbar_init(&bars[EW_KQ], 1);
        fence_mbar_init();
    }
    if (warp == 8) tmem_alloc<256>(tmem_slot);
    if (warp == 9 && lane == 0) {
        tma_prefetch_desc(&tmQKV);
        tma_prefetch_desc(&tmCTX);
    }
    __syncthreads();

    // ---- stage Q and K of the head: 64-row boxes of 64 bytes through TMA (the 64-byte swizzle is the layout the
    //      MMA descriptors expect).  Thread-per-row loads of 64-byte rows 2304 bytes apart cost 32 LSU wavefronts
    //      per instruction, and ncu showed the LSU data pipe -- not the SFU, not the tensor pipe -- as this
    //      kernel's busiest unit (58 % of peak over the whole kernel) ----
    if (warp == 9) {
        if (lane == 0) {
            mbar_expect_tx(&bars[EW_KQ], 2u * (uint32_t)NB * 4096u);
            for (int i = 0; i < NB; i++) {
                tma_load_2d(smem + EW_Q_OFF + i * 4096, &tmQKV, h * EA_DH, (int)(row0 + i * 64), &bars[EW_KQ]);
                tma_load_2d(smem + EW_K_OFF + i * 4096, &tmQKV, p.H + h * EA_DH, (int)(row0 + i * 64), &bars[EW_KQ]);
            }
        }
        // the ones operand of the row sums: [16 rows x 128 B], row 0 = 1.0 (a row is 128 contiguous bytes whatever
        // the swizzle does inside it)
        for (int i = lane; i < 128; i += 32) {
            const uint32_t w = i < 8 ? 0x3f803f80u : 0u;
            *reinterpret_cast<uint4*>(smem + EW_ONE_OFF + i * 16) = make_uint4(w, w, w, w);
        }
    } else if (warp < 8) {
        const bool in = tid < T;  // T is a multiple of 64: warp-uniform
        const float bias = in ? p.mask_bias[row0 + tid] : -1e30f;
        const uint32_t keep = __ballot_sync(0xffffffffu, bias == 0.f);
        if (lane == 0) keep_s[warp] = keep;
        if (!in && tid < QT * 128)  // query rows past T inside the last tile are zero
            for (int c = 0; c < 4; c++) *reinterpret_cast<uint4*>(smem + EW_Q_OFF + sw64_off(tid, c)) = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async_smem();
    mbar_wait(&bars[EW_KQ], 0);
    // ||q||^2 per row and the largest ||k||^2 of the head: every score obeys s <= ||q|| max||k|| (Cauchy-Schwarz)
    auto sumsq_row = [&](const unsigned char* tile, int r) {
        float acc = 0.f;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint4 w = *reinterpret_cast<const uint4*>(tile + sw64_off(r, c));
            const uint32_t e[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float lo = __uint_as_float(e[i] << 16), hi = __uint_as_float(e[i] & 0xffff0000u);
                acc = fmaf(lo, lo, acc);
                acc = fmaf(hi, hi, acc);
            }
        }
        return acc;
    };
    float qn2 = 0.f;
    if (warp < 8) {
        float kn = 0.f;
        if (tid < T) {
            qn2 = sumsq_row(smem + EW_Q_OFF, tid);
            kn = sumsq_row(smem + EW_K_OFF, tid);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) kn = fmaxf(kn, __shfl_xor_sync(0xffffffffu, kn, o));
        if (lane == 0) kmax_s[warp] = kn;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const float kn2 = fmaxf(fmaxf(fmaxf(kmax_s[0], kmax_s[1]), fmaxf(kmax_s[2], kmax_s[3])),
                            fmaxf(fmaxf(kmax_s[4], kmax_s[5]), fmaxf(kmax_s[6], kmax_s[7])));
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

    // which 64-key blocks take part, and which of them have no masked key (CTA-uniform)
    uint32_t act = 0, full = 0;
    for (int j = 0; j < NB; j++) {
        const uint32_t w0 = keep_s[2 * j], w1 = keep_s[2 * j + 1];
        if (w0 | w1) act |= 1u << j;
        if ((w0 & w1) == 0xffffffffu) full |= 1u << j;
    }
    // A chunk whose every key is masked: all its scores carry the same -1e30 bias, which swallows them in fp32, and
    // the softmax is uniform over all T keys.  Scale 0 and shift 0 over all blocks give exactly that.
    const bool no_key = act == 0;
    if (no_key) act = full = (1u << NB) - 1u;
    const float scale = no_key ? 0.f : p.scale_log2;

    // the softmax shift: see enc_attention_blk_kernel
    const int g = warp >> 2, row = tid & 127;  // softmax roles (warps 0-7)
    const float mhat = sqrtf(qn2 * kn2) * scale * 1.0001f + 1e-3f;
    const bool exact_max = __syncthreads_or(mhat > p.shift_limit ? 1 : 0) != 0 && !no_key;
    const uint32_t lane_addr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(g & 1) * 128u;
    float mx = mhat;
    if (exact_max) {
        // ---- pass 1 (only when the bound is not usable): the exact row maximum over the kept keys, 128 keys at a
        //      time, both tiles at once into their groups' columns ----
        mx = -INFINITY;
        uint32_t ph = 0;
        uint32_t r[32];
        for (int k0 = 0; k0 < T; k0 += 128) {
            const int n1 = min(128, T - k0);
            if (warp == 8) {
                if (elect_one()) {
                    const uint32_t idesc = make_idesc_bf16(128, n1);
                    for (int gg = 0; gg < QT; gg++) {
                        const uint64_t dq = make_smem_desc_sw64(smem_u32(smem + EW_Q_OFF + gg * 8192));
                        const uint64_t dk = make_smem_desc_sw64(smem_u32(smem + EW_K_OFF + k0 * 64));
#pragma unroll
                        for (int kk = 0; kk < EA_DH / 16; kk++)
                            umma_bf16(tmem_base + gg * 128, dq + 2 * kk, dk + 2 * kk, idesc, kk != 0 ? 1u : 0u);
                    }
                    umma_commit(&bars[EW_P1]);
                }
                __syncwarp();
            }
            mbar_wait(&bars[EW_P1], ph);
            ph ^= 1u;
            tc_fence_after();
            if (warp < 8 && g < QT) {
                for (int c = 0; c < n1; c += 32) {
                    tmem_ld_32x32(lane_addr + c, r);
                    tmem_ld_wait();
                    const uint32_t w = keep_s[(k0 + c) >> 5];
#pragma unroll
                    for (int j = 0; j < 32; j++)
                        if ((w >> j) & 1u) mx = fmaxf(mx, __uint_as_float(r[j]) * scale);
                }
            }
            tc_fence_before();
            __syncthreads();  // every row has read this piece before the next MMA overwrites the columns
            tc_fence_after();
        }
    }

    const int j_first = __ffs(act) - 1;
    const int nact = __popc(act);
    if (warp < 8) {
        // ===================== softmax groups =====================
        if (g < QT) {
            const float neg_mx = -mx;
            unsigned char* Pb = smem + EW_P_OFF + g * 16384;
            uint32_t r[2][32];
            int idx = 0;
            for (int j = 0; j < NB; j++) {
                if (!((act >> j) & 1u)) continue;
                mbar_wait(&bars[EW_S_FULL + g], idx & 1);
                tc_fence_after();
                tmem_ld_32x32(lane_addr, r[0]);
                tmem_ld_32x32(lane_addr + 32, r[1]);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(&bars[EW_S_FREE + g]);  // the next block's scores may overwrite the columns
                const bool unmasked = (full >> j) & 1u;
                uint32_t pk[32];
#pragma unroll
                for (int hh = 0; hh < 2; hh++) {
                    if (unmasked) {
#pragma unroll
                        for (int jj = 0; jj < 16; jj++) {
                            const float x0 = fmaf(__uint_as_float(r[hh][2 * jj]), scale, neg_mx);
                            const float x1 = fmaf(__uint_as_float(r[hh][2 * jj + 1]), scale, neg_mx);
                            pk[hh * 16 + jj] = pack_bf16x2_rn(ex2_fast(x0), ex2_fast(x1));
                        }
                    } else {
                        const uint32_t w = keep_s[2 * j + hh];
#pragma unroll
                        for (int jj = 0; jj < 16; jj++) {
                            const float p0 = (w >> (2 * jj)) & 1u ? ex2_fast(fmaf(__uint_as_float(r[hh][2 * jj]), scale, neg_mx)) : 0.f;
                            const float p1 = (w >> (2 * jj + 1)) & 1u ? ex2_fast(fmaf(__uint_as_float(r[hh][2 * jj + 1]), scale, neg_mx)) : 0.f;
                            pk[hh * 16 + jj] = pack_bf16x2_rn(p0, p1);
                        }
                    }
                }
                // the previous block's product has read P
                if (idx > 0) mbar_wait(&bars[EW_PV_DONE + g], (idx - 1) & 1);
#pragma unroll
                for (int i = 0; i < 8; i++)
                    *reinterpret_cast<uint4*>(Pb + sw128_off(row, i)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
                fence_proxy_async_smem();
                mbar_arrive(&bars[EW_P_FULL + g]);
                idx++;
            }
            mbar_wait(&bars[EW_PV_DONE + g], (nact - 1) & 1);
            tc_fence_after();
            // ---- normalise and store: columns 64..95 = O, column 96 = the row sum ----
            uint32_t rs[16];
            tmem_ld_32x32(lane_addr + 64, r[0]);
            tmem_ld_32x16(lane_addr + 96, rs);
            tmem_ld_wait();
            {
                // the context rows leave through the group's P buffer (free now) and one TMA store per 64 rows
                const float inv = 1.0f / __uint_as_float(rs[0]);
                uint32_t o[16];
#pragma unroll
                for (int jj = 0; jj < 16; jj++)
                    o[jj] = pack_bf16x2_rn(__uint_as_float(r[0][2 * jj]) * inv, __uint_as_float(r[0][2 * jj + 1]) * inv);
#pragma unroll
                for (int i = 0; i < 4; i++)
                    *reinterpret_cast<uint4*>(Pb + sw64_off(row, i)) = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
                fence_proxy_async_smem();
                asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
                if (row == 0) {
                    for (int hf = 0; hf < 2; hf++)
                        if (g * 128 + hf * 64 < T)
                            tma_store_2d(&tmCTX, h * EA_DH, (int)(row0 + g * 128 + hf * 64), Pb + hf * 4096);
                    bulk_commit_group();
                    bulk_wait_group_read0();
                }
            }
        }
    } else if (warp == 8) {
        // ===================== MMA issuer (converged warp, one elected lane issues) =====================
        const uint32_t idesc_s = make_idesc_bf16(128, 64);
        const uint32_t idesc_o = make_idesc_bf16(128, EA_DH) | (1u << 16);  // B = V is MN-major: [token][dim]
        const uint32_t idesc_1 = make_idesc_bf16(128, 16);
        const uint64_t d_one = make_smem_desc_sw128(smem_u32(smem + EW_ONE_OFF));
        auto issue_scores = [&](int gg, int j) {
            if (elect_one()) {
                const uint64_t dq = make_smem_desc_sw64(smem_u32(smem + EW_Q_OFF + gg * 8192));
                const uint64_t dk = make_smem_desc_sw64(smem_u32(smem + EW_K_OFF + j * 64 * 64));
#pragma unroll
                for (int kk = 0; kk < EA_DH / 16; kk++)
                    umma_bf16(tmem_base + gg * 128, dq + 2 * kk, dk + 2 * kk, idesc_s, kk != 0 ? 1u : 0u);
                umma_commit(&bars[EW_S_FULL + gg]);
            }
            __syncwarp();
        };
        for (int gg = 0; gg < QT; gg++) issue_scores(gg, j_first);
        int idx = 0;
        for (int j = 0; j < NB; j++) {
            if (!((act >> j) & 1u)) continue;
            const int slot = idx & 1, use = idx >> 1;
            const uint32_t rest = act >> (j + 1);
            if (rest) {
                const int jn = j + __ffs(rest);
                for (int gg = 0; gg < QT; gg++) {
                    mbar_wait(&bars[EW_S_FREE + gg], idx & 1);
                    tc_fence_after();
                    issue_scores(gg, jn);
                }
            }
            mbar_wait(&bars[EW_V_FULL + slot], use & 1);
            for (int gg = 0; gg < QT; gg++) {
                mbar_wait(&bars[EW_P_FULL + gg], idx & 1);
                tc_fence_after();
                if (elect_one()) {
                    const uint64_t dp = make_smem_desc_sw128(smem_u32(smem + EW_P_OFF + gg * 16384));
                    const uint64_t dv = make_smem_desc_sw64(smem_u32(smem + EW_V_OFF + slot * 4096));
#pragma unroll
                    for (int kk = 0; kk < 4; kk++) {  // 16 keys per step: 32 bytes along a P row, two 8-token groups of V
                        umma_bf16(tmem_base + gg * 128 + 64, dp + 2 * kk, dv + 64 * kk, idesc_o, (idx | kk) != 0 ? 1u : 0u);
                        umma_bf16(tmem_base + gg * 128 + 96, dp + 2 * kk, d_one + 2 * kk, idesc_1, (idx | kk) != 0 ? 1u : 0u);
                    }
                    umma_commit(&bars[EW_PV_DONE + gg]);
                }
                __syncwarp();
            }
            if (elect_one()) umma_commit(&bars[EW_V_FREE + slot]);  // both groups' products have read the slot
            __syncwarp();
            idx++;
        }
        // nobody waits for the last commit on either slot: see them land before the CTA gives its shared memory back
        for (int k = nact - 1; k >= 0 && k >= nact - 2; k--) mbar_wait(&bars[EW_V_FREE + (k & 1)], (k >> 1) & 1);
    } else {
        // ===================== V loader: one 64-token block per TMA box, two slots =====================
        if (lane == 0) {
            int idx = 0;
            for (int j = 0; j < NB; j++) {
                if (!((act >> j) & 1u)) continue;
                const int slot = idx & 1, use = idx >> 1;
                if (use > 0) mbar_wait(&bars[EW_V_FREE + slot], (use - 1) & 1);
                mbar_expect_tx(&bars[EW_V_FULL + slot], 4096u);
                tma_load_2d(smem + EW_V_OFF + slot * 4096, &tmQKV, 2 * p.H + h * EA_DH, (int)(row0 + j * 64), &bars[EW_V_FULL + slot]);
                idx++;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

// ---- E3d: the same roles in persistent CTAs ----------------------------------------------------------------
// enc_attention_ws_kernel's timeline (clock64 stamps of one row): of the 18.3 k cycles a CTA lives, 5.4 k pass
// before the first block (barrier init, TMEM allocation, the Q/K round trip, the norms behind two CTA-wide
// barriers) and 1.7 k after the last.  Here two CTAs per SM stay resident and walk over the (head, chunk) items:
//   * Q and K of item i+1 land in a second buffer while item i is computed; warp 9 also derives the item's
//     "meta" from them ahead of time -- which key blocks take part, which have no masked key, and the softmax
//     shift.  The shift is the same for every row of the item: sqrt(max||q||^2 max||k||^2) bounds every score of
//     the item, and a common shift is as good as a per-row one while nothing underflows (see the blocked kernel);
//   * the MMA warp issues the first scores of item i+1 as soon as the rows have read the last scores of item i,
//     so a row goes from its last block straight into the epilogue and from there into ready scores;
//   * barriers are initialised and TMEM is allocated once per CTA; phases run on across items.
// An item whose bound is too large for the shift takes the exact row maxima in a first sweep over its key
// blocks (scores only) through the same pipeline, then the normal sweep.
constexpr int EP_THREADS = 320;
constexpr int EP_QK_OFF = 0;                // [2 buffers][Q 16 KiB | K 16 KiB], 64-byte swizzle
constexpr int EP_V_OFF = 64 * 1024;         // [2 slots][64 tokens x 64 B]  8 KiB, 64-byte swizzle
constexpr int EP_P_OFF = 72 * 1024;         // [2 groups][128 x 128 B]     32 KiB, 128-byte swizzle (also the output staging)
constexpr int EP_ONE_OFF = 104 * 1024;      // [16 rows x 128 B]            2 KiB: row 0 = ones
constexpr int EP_META_OFF = 106 * 1024;     // [2 buffers][16 words]
constexpr int EP_BAR_OFF = EP_META_OFF + 128;
constexpr int EP_SMEM = EP_BAR_OFF + 256 + 1024 /*alignment*/;
enum { EP_S_FULL = 0, EP_S_FREE = 2, EP_P_FULL = 4, EP_PV_DONE = 6, EP_V_FULL = 8, EP_V_FREE = 10, EP_O_FREE = 12,
       EP_KQ_FULL = 14, EP_KQ_FREE = 16, EP_META_FULL = 18, EP_NBAR = 20 };
// meta words: [0..7] keep bits, [8] act, [9] full, [10] flags (1 = exact maxima, 2 = no key kept), [11] shift (float)

struct EncAttnPersist {
    int T, H, heads;
    int64_t items;      // chunks x heads
    float scale_log2, shift_limit;
    const float* mask_bias;
    __nv_bfloat16* ctx;
};

__global__ void __launch_bounds__(EP_THREADS, 2)
enc_attention_pws_kernel(const __grid_constant__ CUtensorMap tmQKV, const EncAttnPersist p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    unsigned char* smem = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + EP_BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + EP_NBAR);
    volatile uint32_t* meta_all = reinterpret_cast<volatile uint32_t*>(smem + EP_META_OFF);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int T = p.T, NB = T >> 6, QT = (T + 127) >> 7;

    if (tid == 0) {
        for (int g = 0; g < 2; g++) {
            mbar_init(&bars[EP_S_FULL + g], 1);
            mbar_init(&bars[EP_S_FREE + g], 128);
            mbar_init(&bars[EP_P_FULL + g], 128);
            mbar_init(&bars[EP_PV_DONE + g], 1);
            mbar_init(&bars[EP_V_FULL + g], 1);
            mbar_init(&bars[EP_V_FREE + g], 1);
            mbar_init(&bars[EP_O_FREE + g], 128);
            mbar_init(&bars[EP_KQ_FULL + g], 1);
            mbar_init(&bars[EP_KQ_FREE + g], 1);
            mbar_init(&bars[EP_META_FULL + g], 1);
        }
        fence_mbar_init();
    }
    if (warp == 8) tmem_alloc<256>(tmem_slot);
    if (warp == 9) {
        if (lane == 0) tma_prefetch_desc(&tmQKV);
        // the ones operand of the row sums: [16 rows x 128 B], row 0 = 1.0 (a row is 128 contiguous bytes whatever
        // the swizzle does inside it)
        for (int i = lane; i < 128; i += 32) {
            const uint32_t w = i < 8 ? 0x3f803f80u : 0u;
            *reinterpret_cast<uint4*>(smem + EP_ONE_OFF + i * 16) = make_uint4(w, w, w, w);
        }
        // query rows past T inside the last tile stay zero in both buffers (TMA only writes rows < T)
        for (int r = T + lane; r < QT * 128; r += 32)
            for (int bf = 0; bf < 2; bf++)
                for (int c = 0; c < 4; c++) *reinterpret_cast<uint4*>(smem + EP_QK_OFF + bf * 32768 + sw64_off(r, c)) = make_uint4(0, 0, 0, 0);
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
    const int64_t first = blockIdx.x, step = gridDim.x;

    if (warp < 8) {
        // ===================== softmax groups: thread == query row == TMEM lane =====================
        const int g = warp >> 2, row = tid & 127;
        if (g < QT) {
            const uint32_t lane_addr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)g * 128u;
            unsigned char* Pb = smem + EP_P_OFF + g * 16384;
            uint32_t n_it = 0, sjob = 0, pblk = 0;  // items, score jobs and P blocks this group has been through
            uint32_t r[2][32];
            for (int64_t it = first; it < p.items; it += step, n_it++) {
                const int buf = n_it & 1;
                const int h = (int)(it % p.heads);
                const int64_t row0 = (it / p.heads) * T;
                mbar_wait(&bars[EP_META_FULL + buf], (n_it >> 1) & 1);
                const volatile uint32_t* meta = meta_all + buf * 16;
                const uint32_t act = meta[8], full = meta[9], flags = meta[10];
                const bool exact = flags & 1u;
                const float scale = (flags & 2u) ? 0.f : p.scale_log2;
                float mx = __uint_as_float(meta[11]);
                if (exact) {
                    // first sweep: the exact row maximum over the kept keys
                    mx = -INFINITY;
                    for (int j = 0; j < NB; j++) {
                        if (!((act >> j) & 1u)) continue;
                        mbar_wait(&bars[EP_S_FULL + g], sjob & 1);
                        tc_fence_after();
                        tmem_ld_32x32(lane_addr, r[0]);
                        tmem_ld_32x32(lane_addr + 32, r[1]);
                        tmem_ld_wait();
                        tc_fence_before();
                        mbar_arrive(&bars[EP_S_FREE + g]);
                        sjob++;
#pragma unroll
                        for (int hh = 0; hh < 2; hh++) {
                            const uint32_t w = meta[2 * j + hh];
#pragma unroll
                            for (int k = 0; k < 32; k++)
                                if ((w >> k) & 1u) mx = fmaxf(mx, __uint_as_float(r[hh][k]) * scale);
                        }
                    }
                }
                const float neg_mx = -mx;
                bool first_blk = true;
                for (int j = 0; j < NB; j++) {
                    if (!((act >> j) & 1u)) continue;
                    mbar_wait(&bars[EP_S_FULL + g], sjob & 1);
                    tc_fence_after();
                    tmem_ld_32x32(lane_addr, r[0]);
                    tmem_ld_32x32(lane_addr + 32, r[1]);
                    tmem_ld_wait();
                    tc_fence_before();
                    mbar_arrive(&bars[EP_S_FREE + g]);  // the next scores may overwrite the columns
                    sjob++;
                    const bool unmasked = (full >> j) & 1u;
                    uint32_t pk[32];
#pragma unroll
                    for (int hh = 0; hh < 2; hh++) {
                        if (unmasked) {
                            const uint64_t scale2 = f2_pack(scale, scale), nmx2 = f2_pack(neg_mx, neg_mx);
#pragma unroll
                            for (int jj = 0; jj < 16; jj++) {  // s * scale - mx for two keys per FFMA2
                                float x0, x1;
                                f2_unpack(f2_fma(f2_pack_bits(r[hh][2 * jj], r[hh][2 * jj + 1]), scale2, nmx2), x0, x1);
                                pk[hh * 16 + jj] = pack_bf16x2_rn(ex2_fast(x0), ex2_fast(x1));
                            }
                        } else {
                            const uint32_t w = meta[2 * j + hh];
#pragma unroll
                            for (int jj = 0; jj < 16; jj++) {
                                const float p0 = (w >> (2 * jj)) & 1u ? ex2_fast(fmaf(__uint_as_float(r[hh][2 * jj]), scale, neg_mx)) : 0.f;
                                const float p1 = (w >> (2 * jj + 1)) & 1u ? ex2_fast(fmaf(__uint_as_float(r[hh][2 * jj + 1]), scale, neg_mx)) : 0.f;
                                pk[hh * 16 + jj] = pack_bf16x2_rn(p0, p1);
                            }
                        }
                    }
                    // the previous block's product has read P (an item's first block: waited for in the epilogue)
                    if (!first_blk) mbar_wait(&bars[EP_PV_DONE + g], (pblk - 1) & 1);
#pragma unroll
                    for (int i = 0; i < 8; i++)
                        *reinterpret_cast<uint4*>(Pb + sw128_off(row, i)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
                    fence_proxy_async_smem();
                    mbar_arrive(&bars[EP_P_FULL + g]);
                    pblk++;
                    first_blk = false;
                }
                mbar_wait(&bars[EP_PV_DONE + g], (pblk - 1) & 1);
                tc_fence_after();
                // ---- normalise and store: columns 64..95 = O, column 96 = the row sum ----
                uint32_t rs[16];
                tmem_ld_32x32(lane_addr + 64, r[0]);
                tmem_ld_32x16(lane_addr + 96, rs);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(&bars[EP_O_FREE + g]);  // the next item's first product may overwrite O
                const float inv = 1.0f / __uint_as_float(rs[0]);
                uint32_t o[16];
#pragma unroll
                for (int jj = 0; jj < 16; jj++)
                    o[jj] = pack_bf16x2_rn(__uint_as_float(r[0][2 * jj]) * inv, __uint_as_float(r[0][2 * jj + 1]) * inv);
                // through the group's P buffer (free: its last product has completed) so that four lanes store one
                // 64-byte row: thread-per-row stores cost 32 LSU wavefronts per instruction
#pragma unroll
                for (int i = 0; i < 4; i++)
                    *reinterpret_cast<uint4*>(Pb + sw64_off(row, i)) = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
                asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int rr = (warp & 3) * 32 + 8 * i + (lane >> 2), c = lane & 3;
                    const uint4 v = *reinterpret_cast<const uint4*>(Pb + sw64_off(rr, c));
                    const int q = g * 128 + rr;
                    if (q < T) *reinterpret_cast<uint4*>(p.ctx + (size_t)(row0 + q) * p.H + h * EA_DH + c * 8) = v;
                }
                asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");  // staging read before the next P block lands
            }
        }
    } else if (warp == 8) {
        // ===================== MMA issuer (converged warp, one elected lane issues) =====================
        const uint32_t idesc_s = make_idesc_bf16(128, 64);
        const uint32_t idesc_o = make_idesc_bf16(128, EA_DH) | (1u << 16);  // B = V is MN-major: [token][dim]
        const uint32_t idesc_1 = make_idesc_bf16(128, 16);
        const uint64_t d_one = make_smem_desc_sw128(smem_u32(smem + EP_ONE_OFF));
        uint32_t n_it = 0, sjob = 0, pblk = 0, vuse = 0;  // score jobs issued, products issued, V slots consumed
        for (int64_t it = first; it < p.items; it += step, n_it++) {
            const int buf = n_it & 1;
            mbar_wait(&bars[EP_META_FULL + buf], (n_it >> 1) & 1);  // implies Q and K of the item have landed
            const volatile uint32_t* meta = meta_all + buf * 16;
            const uint32_t act = meta[8];
            const bool exact = meta[10] & 1u;
            const int nact = __popc(act);
            const int njobs = exact ? 2 * nact : nact;
            const unsigned char* Qb = smem + EP_QK_OFF + buf * 32768;
            const unsigned char* Kb = Qb + 16384;
            auto issue_scores = [&](int gg, int j, bool last) {
                if (elect_one()) {
                    const uint64_t dq = make_smem_desc_sw64(smem_u32(Qb + gg * 8192));
                    const uint64_t dk = make_smem_desc_sw64(smem_u32(Kb + j * 64 * 64));
#pragma unroll
                    for (int kk = 0; kk < EA_DH / 16; kk++)
                        umma_bf16(tmem_base + gg * 128, dq + 2 * kk, dk + 2 * kk, idesc_s, kk != 0 ? 1u : 0u);
                    umma_commit(&bars[EP_S_FULL + gg]);
                    if (last && gg == QT - 1) umma_commit(&bars[EP_KQ_FREE + buf]);  // Q and K of the item have been read
                }
                __syncwarp();
            };
            // job k of the item reads key block blk(k): the k-th set bit of act, twice over when exact
            auto blk_of = [&](int k) {
                int kk = k >= nact ? k - nact : k;
                uint32_t a = act;
                for (int i = 0; i < kk; i++) a &= a - 1;
                return __ffs(a) - 1;
            };
            // the item's first scores: the rows have read the previous item's last ones
            for (int gg = 0; gg < QT; gg++) {
                if (sjob > 0) {
                    mbar_wait(&bars[EP_S_FREE + gg], (sjob - 1) & 1);
                    tc_fence_after();
                }
                issue_scores(gg, blk_of(0), njobs == 1);
            }
            for (int k = 0; k < njobs; k++, sjob++) {
                if (k + 1 < njobs) {
                    const int jn = blk_of(k + 1);
                    for (int gg = 0; gg < QT; gg++) {
                        mbar_wait(&bars[EP_S_FREE + gg], sjob & 1);
                        tc_fence_after();
                        issue_scores(gg, jn, k + 2 == njobs);
                    }
                }
                if (exact && k < nact) continue;  // first sweep: scores only
                const bool first_blk = exact ? k == nact : k == 0;
                const int slot = vuse & 1;
                mbar_wait(&bars[EP_V_FULL + slot], (vuse >> 1) & 1);
                for (int gg = 0; gg < QT; gg++) {
                    mbar_wait(&bars[EP_P_FULL + gg], pblk & 1);
                    if (first_blk && n_it > 0) mbar_wait(&bars[EP_O_FREE + gg], (n_it - 1) & 1);  // the rows have read the previous O
                    tc_fence_after();
                    if (elect_one()) {
                        const uint64_t dp = make_smem_desc_sw128(smem_u32(smem + EP_P_OFF + gg * 16384));
                        const uint64_t dv = make_smem_desc_sw64(smem_u32(smem + EP_V_OFF + slot * 4096));
#pragma unroll
                        for (int kk = 0; kk < 4; kk++) {  // 16 keys per step: 32 bytes along a P row, two 8-token groups of V
                            umma_bf16(tmem_base + gg * 128 + 64, dp + 2 * kk, dv + 64 * kk, idesc_o, (!first_blk || kk != 0) ? 1u : 0u);
                            umma_bf16(tmem_base + gg * 128 + 96, dp + 2 * kk, d_one + 2 * kk, idesc_1, (!first_blk || kk != 0) ? 1u : 0u);
                        }
                        umma_commit(&bars[EP_PV_DONE + gg]);
                    }
                    __syncwarp();
                }
                if (elect_one()) umma_commit(&bars[EP_V_FREE + slot]);  // both groups' products have read the slot
                __syncwarp();
                pblk++;
                vuse++;
            }
        }
        // nobody waits for the last commits on the V slots and the Q/K buffers: see them land before the CTA gives
        // its shared memory back
        for (uint32_t k = vuse; k > 0 && k + 2 > vuse; k--) mbar_wait(&bars[EP_V_FREE + ((k - 1) & 1)], ((k - 1) >> 1) & 1);
        for (uint32_t k = n_it; k > 0 && k + 2 > n_it; k--) mbar_wait(&bars[EP_KQ_FREE + ((k - 1) & 1)], ((k - 1) >> 1) & 1);
    } else {
        // ===================== loader: Q / K of the next item, its meta, V blocks of the current one =====================
        uint32_t n_it = 0, vuse = 0;
        auto load_qk = [&](int64_t it, uint32_t n) {  // lane 0
            const int buf = n & 1;
            if (n >= 2) mbar_wait(&bars[EP_KQ_FREE + buf], ((n >> 1) - 1) & 1);
            const int h = (int)(it % p.heads);
            const int64_t row0 = (it / p.heads) * T;
            unsigned char* Qb = smem + EP_QK_OFF + buf * 32768;
            mbar_expect_tx(&bars[EP_KQ_FULL + buf], 2u * (uint32_t)NB * 4096u);
            for (int i = 0; i < NB; i++) {
                tma_load_2d(Qb + i * 4096, &tmQKV, h * EA_DH, (int)(row0 + i * 64), &bars[EP_KQ_FULL + buf]);
                tma_load_2d(Qb + 16384 + i * 4096, &tmQKV, p.H + h * EA_DH, (int)(row0 + i * 64), &bars[EP_KQ_FULL + buf]);
            }
        };
        auto sumsq_row = [&](const unsigned char* tile, int r) {
            float acc = 0.f;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const uint4 w = *reinterpret_cast<const uint4*>(tile + sw64_off(r, c));
                const uint32_t e[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const float lo = __uint_as_float(e[i] << 16), hi = __uint_as_float(e[i] & 0xffff0000u);
                    acc = fmaf(lo, lo, acc);
                    acc = fmaf(hi, hi, acc);
                }
            }
            return acc;
        };
        // whole warp.  The buffer's previous tenant is item n - 2; `prev_use` = V slots consumed up to and including
        // that item: once its last product has completed, no row and no MMA looks at its meta any more.
        auto make_meta = [&](int64_t it, uint32_t n, uint32_t prev_use) {
            const int buf = n & 1;
            if (prev_use > 0) mbar_wait(&bars[EP_V_FREE + ((prev_use - 1) & 1)], ((prev_use - 1) >> 1) & 1);
            const int64_t row0 = (it / p.heads) * T;
            volatile uint32_t* meta = meta_all + buf * 16;
            uint32_t act = 0, full = 0;
            for (int w = 0; w < 2 * NB; w++) {
                const uint32_t keep = __ballot_sync(0xffffffffu, p.mask_bias[row0 + w * 32 + lane] == 0.f);
                if (lane == 0) meta[w] = keep;
                if (keep) act |= 1u << (w >> 1);
                if (keep != 0xffffffffu) full |= 1u << (w >> 1);  // collects the blocks that are NOT full
            }
            full = ~full & ((1u << NB) - 1u);
            const bool no_key = act == 0;
            if (no_key) act = full = (1u << NB) - 1u;
            mbar_wait(&bars[EP_KQ_FULL + buf], (n >> 1) & 1);
            const unsigned char* Qb = smem + EP_QK_OFF + buf * 32768;
            float qm = 0.f, km = 0.f;
            for (int r = lane; r < T; r += 32) {
                qm = fmaxf(qm, sumsq_row(Qb, r));
                km = fmaxf(km, sumsq_row(Qb + 16384, r));
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                qm = fmaxf(qm, __shfl_xor_sync(0xffffffffu, qm, o));
                km = fmaxf(km, __shfl_xor_sync(0xffffffffu, km, o));
            }
            // every score of the item obeys |s| <= max||q|| max||k|| (Cauchy-Schwarz)
            const float mhat = no_key ? 1e-3f : sqrtf(qm * km) * p.scale_log2 * 1.0001f + 1e-3f;
            const bool exact = mhat > p.shift_limit && !no_key;
            if (lane == 0) {
                meta[8] = act;
                meta[9] = full;
                meta[10] = (exact ? 1u : 0u) | (no_key ? 2u : 0u);
                meta[11] = __float_as_uint(mhat);
                __threadfence_block();
                mbar_arrive(&bars[EP_META_FULL + buf]);
            }
            __syncwarp();
        };
        if (first < p.items) {
            if (lane == 0) load_qk(first, 0);
            __syncwarp();
            make_meta(first, 0, 0);
        }
        for (int64_t it = first; it < p.items; it += step, n_it++) {
            const int buf = n_it & 1;
            const int64_t nxt = it + step;
            if (nxt < p.items && lane == 0) load_qk(nxt, n_it + 1);
            __syncwarp();
            const int h = (int)(it % p.heads);
            const int64_t row0 = (it / p.heads) * T;
            const uint32_t act = meta_all[buf * 16 + 8];
            const uint32_t use0 = vuse;  // V slots consumed by the items before this one
            int idx = 0;
            bool meta_done = nxt >= p.items;
            for (int j = 0; j < NB; j++) {
                if (!((act >> j) & 1u)) continue;
                if (idx == 2 && !meta_done) {  // two V blocks are under way: time for the next item's meta
                    make_meta(nxt, n_it + 1, 0);  // (the second block's slot wait has already seen the previous item out)
                    meta_done = true;
                }
                if (lane == 0) {
                    const int slot = vuse & 1;
                    if (vuse >= 2) mbar_wait(&bars[EP_V_FREE + slot], ((vuse >> 1) - 1) & 1);
                    mbar_expect_tx(&bars[EP_V_FULL + slot], 4096u);
                    tma_load_2d(smem + EP_V_OFF + slot * 4096, &tmQKV, 2 * p.H + h * EA_DH, (int)(row0 + j * 64), &bars[EP_V_FULL + slot]);
                }
                __syncwarp();
                vuse++;
                idx++;
            }
            // with a single block issued, nothing has waited for the previous item's last product yet -- and no new use
            // of that product's slot exists, so waiting for its phase here cannot be overtaken by the next one
            if (!meta_done) make_meta(nxt, n_it + 1, idx >= 2 ? 0 : use0);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

int launch_enc_attention(const __nv_bfloat16* qkv, const float* mask_bias, __nv_bfloat16* ctx, int64_t B, int T, int H,
                         int heads, cudaStream_t st) {
    if (B <= 0) return 0;
    if (H != heads * EA_DH) return fail("encoder attention: head dimension must be 32");
    if (T <= 0 || T > EA_MAXT || (T & 63)) return fail("encoder attention: sequence length must be a multiple of 64, <= 256");
    if (B > 65535) return fail("encoder attention: more than 65535 chunks per launch");
    static bool attr_set[64] = {false};
    static const int variant = [] {  // BFK_ENC_ATTN=row: the first version (whole score row in TMEM); =blk: the second
        const char* e = getenv("BFK_ENC_ATTN");
        return e && e[0] == 'r' ? 0 : e && e[0] == 'b' ? 1 : e && e[0] == 'w' ? 2 : 3;
    }();
    const bool whole_row = variant == 0;
    int dev = 0;
    BFK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        BFK_CUDA(cudaFuncSetAttribute(enc_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EA_SMEM));
        BFK_CUDA(cudaFuncSetAttribute(enc_attention_blk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EB_SMEM));
        BFK_CUDA(cudaFuncSetAttribute(enc_attention_ws_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EW_SMEM));
        BFK_CUDA(cudaFuncSetAttribute(enc_attention_pws_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EP_SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    EncAttnParams p;
    p.qkv = qkv;
    p.mask_bias = mask_bias;
    p.ctx = ctx;
    p.T = T;
    p.H = H;
    p.scale_log2 = 1.4426950408889634f / sqrtf((float)EA_DH);
    static const bool exact_shift = [] {  // BFK_ENC_SHIFT=exact: always the true row maximum (tests, comparisons)
        const char* e = getenv("BFK_ENC_SHIFT");
        return e && e[0] == 'e';
    }();
    p.shift_limit = exact_shift ? 0.f : 60.f;
    p.trace = nullptr;
    static const bool trace_on = getenv("BFK_ENC_ATTN_TRACE") != nullptr;
    if (trace_on && variant == 1) {  // debugging (the blocked kernel): synchronous, one CTA's phase timeline on stderr
        unsigned long long* d = nullptr;
        BFK_CUDA(cudaMalloc((void**)&d, 128 * 8));
        BFK_CUDA(cudaMemset(d, 0, 128 * 8));
        p.trace = d;
        dim3 grid((unsigned)heads, (unsigned)B);
        enc_attention_blk_kernel<<<grid, EB_THREADS, EB_SMEM, st>>>(p);
        BFK_CUDA(cudaStreamSynchronize(st));
        unsigned long long hst[128];
        BFK_CUDA(cudaMemcpy(hst, d, sizeof(hst), cudaMemcpyDeviceToHost));
        cudaFree(d);
        fprintf(stderr, "[bfk attn trace] B=%lld T=%d:", (long long)B, T);
        for (int i = 1; i < 128; i++)
            if (hst[i]) fprintf(stderr, " %d:%lld", i, (long long)(hst[i] - hst[0]));
        fprintf(stderr, "\n");
        return 0;
    }
    if (whole_row) {
        dim3 grid((unsigned)((T + 127) / 128), (unsigned)heads, (unsigned)B);
        enc_attention_kernel<<<grid, EA_THREADS, EA_SMEM, st>>>(p);
    } else {
        dim3 grid((unsigned)heads, (unsigned)B);  // one CTA per (head, chunk): K is staged once for its query tiles
        if (variant == 1) enc_attention_blk_kernel<<<grid, EB_THREADS, EB_SMEM, st>>>(p);
        else if (variant == 3) {
            CUtensorMap tmQKV;
            BFK_TRY(make_tmap_head(&tmQKV, qkv, B * T, 3 * H, 64));
            EncAttnPersist pp;
            pp.T = T;
            pp.H = H;
            pp.heads = heads;
            pp.items = B * heads;
            pp.scale_log2 = p.scale_log2;
            pp.shift_limit = p.shift_limit;
            pp.mask_bias = mask_bias;
            pp.ctx = ctx;
            int num_sms = 148;
            cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
            const unsigned ctas = (unsigned)std::min<int64_t>(pp.items, 2 * (int64_t)num_sms);
            enc_attention_pws_kernel<<<ctas, EP_THREADS, EP_SMEM, st>>>(tmQKV, pp);
        } else {
            CUtensorMap tmQKV, tmCTX;
            BFK_TRY(make_tmap_head(&tmQKV, qkv, B * T, 3 * H, 64));
            BFK_TRY(make_tmap_head(&tmCTX, ctx, B * T, H, 64));
            enc_attention_ws_kernel<<<grid, EW_THREADS, EW_SMEM, st>>>(tmQKV, tmCTX, p);
        }
    }
    BFK_CUDA(cudaGetLastError());
    return 0;
}

// ---- E1: embeddings + LayerNorm, and the attention mask as an additive bias ------------------------------
// x[m] = LN(word[ids[m]] + pos[m % T] + type[0]); one warp per token, H = 384 -> 12 values per lane
__global__ void __launch_bounds__(256) enc_embed_ln_kernel(const int64_t* __restrict__ ids, const void* __restrict__ mask,
                                                           int mask_dtype, const float* __restrict__ word,
                                                           const float* __restrict__ pos, const float* __restrict__ type0,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           float eps, int64_t M, int T, int H, int vocab,
                                                           __nv_bfloat16* __restrict__ x, float* __restrict__ mask_bias) {
    const int lane = threadIdx.x & 31;
    const int64_t m = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (m >= M) return;
    int64_t id = ids[m];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    const int t = (int)(m % T);
    float v[16];  // H <= 512
    float s1 = 0.f;
    const int per = (H + 31) / 32;
    const bool vec4 = (H & 127) == 0;  // 128-bit loads: lane owns columns [4*lane + 128*i, +4)
    if (vec4) {
        const float4* w4 = reinterpret_cast<const float4*>(word + (size_t)id * H);
        const float4* p4 = reinterpret_cast<const float4*>(pos + (size_t)t * H);
        const float4* t4 = reinterpret_cast<const float4*>(type0);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (i < (H >> 7)) {
                const float4 a = __ldg(w4 + lane + 32 * i), b = __ldg(p4 + lane + 32 * i), c = __ldg(t4 + lane + 32 * i);
                v[4 * i + 0] = a.x + b.x + c.x;
                v[4 * i + 1] = a.y + b.y + c.y;
                v[4 * i + 2] = a.z + b.z + c.z;
                v[4 * i + 3] = a.w + b.w + c.w;
                s1 += (v[4 * i] + v[4 * i + 1]) + (v[4 * i + 2] + v[4 * i + 3]);
            }
        }
    } else {
        for (int i = 0; i < per; i++) {
            const int c = lane + 32 * i;
            v[i] = c < H ? word[(size_t)id * H + c] + pos[(size_t)t * H + c] + type0[c] : 0.f;
            s1 += v[i];
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    const float mean = s1 / H;
    float s2 = 0.f;
    if (vec4) {
#pragma unroll
        for (int i = 0; i < 16; i++)
            if (i < (H >> 5)) {
                const float dlt = v[i] - mean;
                s2 = fmaf(dlt, dlt, s2);
            }
    } else {
        for (int i = 0; i < per; i++) {
            const int c = lane + 32 * i;
            const float dlt = c < H ? v[i] - mean : 0.f;
            s2 = fmaf(dlt, dlt, s2);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    const float rstd = rsqrtf(s2 / H + eps);
    if (vec4) {
        const float4* g4 = reinterpret_cast<const float4*>(gamma);
        const float4* b4 = reinterpret_cast<const float4*>(beta);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (i < (H >> 7)) {
                const float4 g = __ldg(g4 + lane + 32 * i), b = __ldg(b4 + lane + 32 * i);
                const __nv_bfloat162 lo = __floats2bfloat162_rn((v[4 * i] - mean) * rstd * g.x + b.x, (v[4 * i + 1] - mean) * rstd * g.y + b.y);
                const __nv_bfloat162 hi = __floats2bfloat162_rn((v[4 * i + 2] - mean) * rstd * g.z + b.z, (v[4 * i + 3] - mean) * rstd * g.w + b.w);
                uint2 o;
                o.x = *reinterpret_cast<const uint32_t*>(&lo);
                o.y = *reinterpret_cast<const uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(x + (size_t)m * H + 4 * (lane + 32 * i)) = o;
            }
        }
    } else {
        for (int i = 0; i < per; i++) {
            const int c = lane + 32 * i;
            if (c < H) x[(size_t)m * H + c] = __float2bfloat16_rn((v[i] - mean) * rstd * gamma[c] + beta[c]);
        }
    }
    if (lane == 0) {
        float on;
        switch (mask_dtype) {
            case BFK_DT_I64: on = (float)reinterpret_cast<const long long*>(mask)[m]; break;
            case BFK_DT_I32: on = (float)reinterpret_cast<const int*>(mask)[m]; break;
            default: on = (float)reinterpret_cast<const unsigned char*>(mask)[m]; break;
        }
        mask_bias[m] = on != 0.f ? 0.f : -1e30f;
    }
}

int launch_enc_embed(const int64_t* ids, const void* mask, int mask_dtype, const float* word, const float* pos,
                     const float* type0, const float* gamma, const float* beta, float eps, int64_t M, int T, int H,
                     int vocab, __nv_bfloat16* x, float* mask_bias, cudaStream_t st) {
    if (M <= 0) return 0;
    if (H > 512) return fail("encoder: hidden size > 512 is not supported");
    const int64_t blocks = (M * 32 + 255) / 256;
    enc_embed_ln_kernel<<<(unsigned)blocks, 256, 0, st>>>(ids, mask, mask_dtype, word, pos, type0, gamma, beta, eps, M, T, H,
                                                         vocab, x, mask_bias);
    BFK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace bfk

// Shared helpers for libbfk.so (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/bfk.h"

namespace bfk {

// ---- error plumbing -------------------------------------------------------
void set_error(const std::string& msg);
int fail(const std::string& msg);  // sets the message, returns 1

#define BFK_CUDA(expr)                                                                      \
    do {                                                                                    \
        cudaError_t _e = (expr);                                                            \
        if (_e != cudaSuccess) {                                                            \
            return ::bfk::fail(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + \
                               __FILE__ + ":" + std::to_string(__LINE__) + ")");           \
        }                                                                                   \
    } while (0)

#define BFK_TRY(expr)          \
    do {                       \
        int _rc = (expr);      \
        if (_rc) return _rc;   \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = (cudaSetDevice(dev) == cudaSuccess);
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

static inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
static inline int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

// ---- the index object -----------------------------------------------------
// Data layout in HBM (DESIGN.md "Data layout"):
//   x32  [cap, ld32] fp32 master rows, ld32 = round_up(d, 4)   (BFK_STORE_F32 only)
//   x16  [cap, ld16] bf16 scan copy,   ld16 = round_up(d, 64)  (zero padded)
//   nrm  [cap]       fp32 squared norms of the stored (fp32 or bf16) rows
//   maxn [2]         fp32 bits of max_i ||x_i||^2 and of max_i ||x_i - bf16(x_i)||^2 (monotone under atomicMax)
struct Workspace {
    void* ptr = nullptr;
    size_t bytes = 0;
};

// a device array that grows in place (bfk_vmem.cu): address range reserved once, physical memory mapped in chunks
struct VmArray;
int vm_create(int device, size_t va_bytes, VmArray** out);  // non-zero: virtual memory management unavailable
int vm_grow(VmArray* a, size_t bytes);                       // non-zero: out of device memory (array unchanged)
void* vm_ptr(const VmArray* a);
size_t vm_mapped(const VmArray* a);
void vm_destroy(VmArray* a);

}  // namespace bfk

struct bfk_index {
    int d = 0;
    int metric = BFK_METRIC_L2;
    int device = 0;
    int storage = BFK_STORE_F32;
    int ld32 = 0;
    int ld16 = 0;
    int num_sms = 148;
    int64_t ntotal = 0;
    int64_t cap = 0;
    float* x32 = nullptr;
    __nv_bfloat16* x16 = nullptr;
    float* nrm = nullptr;
    // the three arrays above live in growable virtual-memory arrays when the driver offers them (growth maps more
    // physical memory behind the same addresses: no copy, no second resident copy); all null -> cudaMalloc + copy
    bfk::VmArray* vm32 = nullptr;
    bfk::VmArray* vm16 = nullptr;
    bfk::VmArray* vmn = nullptr;
    unsigned int* maxn = nullptr;
    bfk::Workspace ws;       // search scratch, grow-only
    cudaEvent_t ws_free = nullptr;  // recorded (lazily, see ws_settle) behind the last work that used `ws`
    cudaStream_t ws_stream = nullptr;  // ... and the stream it ran on
    bool ws_stream_known = false;
    bool ws_dirty = false;          // work using `ws` was enqueued on ws_stream and no event has been recorded behind it yet
    int* h_flag = nullptr;          // pinned host words: [0] count of uncertified queries of the last tensor pass,
                                    // [1] that pass's sequence number, written after the count (the host polls it)
    int flag_seq = 0;               // sequence number of the last tensor pass / host-notifying search enqueued
    int* sel_ticket = nullptr;      // device word for K4's "last CTA" ticket (SelectNotify)
    int* h_flag_dev = nullptr;      // the same word as the GPU addresses it (K5's last CTA writes it)
    char* h_stage = nullptr;        // pinned staging for small host-side searches (queries in, (D, I) out)
    // id selector of the search in progress (bfk_search_sel; null otherwise), set and cleared under `mu`:
    //   sel       bit i set <=> row i may be returned (faiss IDSelectorBitmap layout)
    //   sel_bias  [tiles*256] fp32: the tensor scan's per-row bias with +inf for rows that are not selected
    const unsigned char* sel = nullptr;
    const float* sel_bias = nullptr;
    void* sel_buf = nullptr;        // grow-only device buffer behind sel / sel_bias
    size_t sel_buf_bytes = 0;
    // Appends (K1 / K2) are only enqueued on the caller's stream.  `rows_ready` is recorded behind each of them;
    // every path that reads the rows on another stream (regrowth copies, reconstruct / write_index, remove_ids,
    // a search on a different stream) first makes its stream wait for it.
    cudaEvent_t rows_ready = nullptr;
    bool rows_pending = false;      // an append has been enqueued since the last full synchronisation
    cudaStream_t rows_stream = nullptr;  // the stream of that append
    cudaEvent_t stage_done = nullptr;    // the last H2D copy out of `h_stage` has completed
    // tensor maps of the scan, re-encoded only when what they describe changes
    //   [0] corpus scan copy: key (x16, ntotal, box rows)   [1] padded bf16 queries: key (q16, nq_pad)
    struct TmapSlot {
        alignas(64) unsigned char map[128];
        const void* base = nullptr;
        int64_t rows = -1;
        int box_rows = 0;
    } tmap[2];
    // bfk_profile_begin/end: device timestamps of every tensor-scan launch, [slot][0] = first CTA in,
    // [slot][1] = last CTA out (globaltimer ns), collected without synchronising inside the searches
    unsigned long long* prof_ring = nullptr;
    int64_t prof_cap = 0;
    int64_t prof_used = 0;
    bool prof_on = false;
    mutable std::mutex mu;
};

namespace bfk {

int ensure_capacity(bfk_index* idx, int64_t need, cudaStream_t st);
// grows the storage (amortised doubling) under the index mutex; rows appended so far are carried over
int grow_for(bfk_index* idx, int64_t need, cudaStream_t st);
// makes `st` wait for the appends enqueued so far (no-op when none is pending)
int wait_rows_ready(const bfk_index* idx, cudaStream_t st);
// records `rows_ready` behind an append just enqueued on `st`
int mark_rows_ready(bfk_index* idx, cudaStream_t st);
int ensure_workspace(bfk_index* idx, size_t bytes);

// kernels / launchers implemented across the .cu files -----------------------
// K2: append rows (cast + norms)
int launch_append_f32(bfk_index* idx, const float* src_dev, int64_t src_ld, int64_t n, cudaStream_t st);
// K1: pool + normalise (+ optional append into the index tail)
int launch_pool(const void* tok, int tok_dtype, const void* mask, int mask_dtype, int64_t B, int T,
                int H, float* out32, int64_t out32_ld, __nv_bfloat16* out16, int64_t out16_ld,
                float* out_nrm, unsigned int* maxn, cudaStream_t st);
// K3a: exact fp32 scan of up to 8 queries; writes sorted partial lists
//   part [nq][nparts][k] composite keys; returns nparts through *nparts_out
int scan_exact_workspace(const bfk_index* idx, int nq, int k, size_t* bytes, int* nparts_out);
//   sel: optional id selector bitmap on the device (bit i <=> row i may be returned), else nullptr
//   fuse: optional -- where (D, I) go when the scan's last CTA may merge the lists itself (small searches);
//         *fused tells whether it will (then no launch_select is needed)
struct ScanFuse {
    float* D;
    int64_t* I;
    int64_t out_ld;
    int64_t id_offset;
    int* ticket;  // device word, 0 between searches
    int* h_seq;   // optional pinned host word that receives `seq` once (D, I) are complete
    int seq;
};
int launch_scan_exact(const bfk_index* idx, const float* q_dev /*[nq, ldq]*/, int ldq, int nq, int k,
                      uint64_t* part, int nparts, cudaStream_t st, const unsigned char* sel,
                      const ScanFuse* fuse = nullptr, bool* fused = nullptr);
// K4/K6: per query, top-k of m composite keys -> (D, I).  `nt` (optional): D / I are pinned host memory and the
// last CTA writes `seq` to the pinned word *h_seq once every query's rows have reached the host
struct SelectNotify {
    int* ticket;  // device word, 0 between searches
    int* h_seq;   // pinned host word as the GPU addresses it
    int seq;
};
int launch_select(const uint64_t* cand, int64_t m_per_query, int64_t nq, int k, int metric,
                  int64_t id_offset, float* D, int64_t* I, int64_t out_ld, cudaStream_t st,
                  const SelectNotify* nt = nullptr);
// waits until the pass / search that carries sequence number idx->flag_seq has published it in idx->h_flag[1]
// (polls the pinned word; falls back to a blocking stream wait for long passes and on errors)
int wait_flag_seq(bfk_index* idx, cudaStream_t st);
// K6 front end: pack (D, I) lists into composite keys
int launch_pack_parts(const float* Dparts, const int64_t* Iparts, int nparts, int64_t nq, int k,
                      int metric, uint64_t* cand, cudaStream_t st);
// selector support: bias[i] = selected(i) ? (l2 ? nrm[i] : 0) : +inf for i < n_pad (rows >= n: +inf)
int launch_sel_bias(const unsigned char* sel, const float* nrm, int64_t n, int64_t n_pad, int is_l2, float* bias,
                    cudaStream_t st);
// K3b + K5 (tensor path); defined in bfk_scan_tc.cu
int tensor_path_supported(const bfk_index* idx, int64_t nq, int k);
size_t tensor_workspace_bytes(const bfk_index* idx, int64_t nq, int k);
int search_tensor(bfk_index* idx, const float* q_dev, int ldq, int64_t nq, int k, float* D_dev, int64_t* I_dev,
                  int64_t id_offset, char* ws_tail, size_t ws_tail_bytes, cudaStream_t st,
                  bfk_search_stats* stats, cudaEvent_t ev0, cudaEvent_t ev1, int defer,
                  const float* q_host = nullptr /* pinned host queries [nq, d]; q_dev then receives a padded copy */,
                  int out_is_host = 0 /* D_dev / I_dev are pinned host memory written by K5 itself */);
int search_tensor_finish(bfk_index* idx, cudaStream_t st, int64_t* n_rewritten);
void search_tensor_forget(bfk_index* idx);
const int* search_tensor_pending_nflag(bfk_index* idx);

// ---- device helpers ---------------------------------------------------------
// order-preserving map float -> uint32 (smaller float <=> smaller uint)
__host__ __device__ __forceinline__ uint32_t f2ord(float f) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; uint32_t u = c.u;
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(uint32_t o) {
    uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
// composite = (ordered key << 32) | row id : total order (key, id)
__host__ __device__ __forceinline__ uint64_t make_comp(float key, uint32_t id) {
    return (uint64_t(f2ord(key)) << 32) | uint64_t(id);
}
constexpr uint64_t COMP_EMPTY = ~0ull;

}  // namespace bfk

// libbfk.so: index object, storage management and the C ABI (include/bfk.h).
//
// Stands in for the faiss.IndexFlat* object that the reference builds in
// semantic-indexer/indexer.py:39,104, fills at :41, persists at :27, reloads at :99 and
// llm-qa/main.py:35, and searches through LangChain at llm-qa/main.py:101,117.
#include <unistd.h>

#include <algorithm>
#include <cfloat>
#include <cstring>
#include <vector>

#include "bfk_common.cuh"

namespace bfk {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
int fail(const std::string& msg) {
    g_err = msg;
    return 1;
}

static int check_device(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail("no CUDA device available: libbfk has no CPU fallback");
    }
    if (device < 0 || device >= n) return fail("device index out of range");
    return 0;
}

// ---- corpus storage: x32 / x16 / nrm ------------------------------------------------------------------
struct StorageSet {
    float* x32 = nullptr;
    __nv_bfloat16* x16 = nullptr;
    float* nrm = nullptr;
    VmArray *vm32 = nullptr, *vm16 = nullptr, *vmn = nullptr;
    int64_t cap = 0;
};

static size_t bytes32(const bfk_index* idx, int64_t rows) { return sizeof(float) * (size_t)rows * idx->ld32; }
static size_t bytes16(const bfk_index* idx, int64_t rows) { return sizeof(__nv_bfloat16) * (size_t)rows * idx->ld16; }
// norms are fetched a whole 256-row tile at a time (1 KiB bulk copies): pad to the tile
static size_t bytesn(int64_t rows) { return sizeof(float) * (size_t)round_up(rows, 256); }

static void free_set(StorageSet* s) {
    if (s->vm16) {
        vm_destroy(s->vm32);
        vm_destroy(s->vm16);
        vm_destroy(s->vmn);
    } else {
        if (s->x32) cudaFree(s->x32);
        if (s->x16) cudaFree(s->x16);
        if (s->nrm) cudaFree(s->nrm);
    }
    *s = StorageSet();
}

// rows the mapped (or allocated) bytes of a virtual-memory set hold
static int64_t vm_rows(const bfk_index* idx, const StorageSet& s) {
    int64_t rows = (int64_t)(vm_mapped(s.vm16) / (sizeof(__nv_bfloat16) * (size_t)idx->ld16));
    if (s.vm32) rows = std::min<int64_t>(rows, (int64_t)(vm_mapped(s.vm32) / (sizeof(float) * (size_t)idx->ld32)));
    int64_t nrows = (int64_t)(vm_mapped(s.vmn) / sizeof(float));
    nrows -= nrows % 256;  // the padded tile must fit as well
    return std::min(rows, nrows);
}

static int vm_set_grow(const bfk_index* idx, StorageSet* s, int64_t rows) {
    if (s->vm32 && vm_grow(s->vm32, bytes32(idx, rows))) return 1;
    if (vm_grow(s->vm16, bytes16(idx, rows))) return 1;
    if (vm_grow(s->vmn, bytesn(rows))) return 1;
    s->cap = vm_rows(idx, *s);
    return 0;
}

// a fresh set with room for `cap` rows: growable arrays when the driver has virtual memory management, else cudaMalloc
static int alloc_set(const bfk_index* idx, int64_t cap, StorageSet* out) {
    *out = StorageSet();
    StorageSet s;
    // address space for as many rows as the GPU's memory could ever hold (a reservation costs nothing)
    size_t free_b = 0, total_b = 0;
    const size_t row_all = bytes16(idx, 1) + 4 + (idx->storage == BFK_STORE_F32 ? bytes32(idx, 1) : 0);
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && total_b > 0) {
        const int64_t max_rows = std::max<int64_t>(std::min<int64_t>(0x7fffffffll, (int64_t)(total_b / row_all)), cap);
        bool ok = vm_create(idx->device, bytes16(idx, max_rows), &s.vm16) == 0 && vm_create(idx->device, bytesn(max_rows), &s.vmn) == 0;
        if (ok && idx->storage == BFK_STORE_F32) ok = vm_create(idx->device, bytes32(idx, max_rows), &s.vm32) == 0;
        if (ok && vm_set_grow(idx, &s, std::max<int64_t>(cap, 1)) == 0) {
            s.x32 = static_cast<float*>(vm_ptr(s.vm32));
            s.x16 = static_cast<__nv_bfloat16*>(vm_ptr(s.vm16));
            s.nrm = static_cast<float*>(vm_ptr(s.vmn));
            *out = s;
            return 0;
        }
        const bool had_vm = s.vm16 != nullptr && s.vmn != nullptr && (idx->storage != BFK_STORE_F32 || s.vm32 != nullptr);
        vm_destroy(s.vm32);
        vm_destroy(s.vm16);
        vm_destroy(s.vmn);
        s = StorageSet();
        if (had_vm) return 1;  // reservations worked, the memory itself ran out
    } else {
        cudaGetLastError();
    }
    cudaError_t e = cudaSuccess;
    if (idx->storage == BFK_STORE_F32) e = cudaMalloc((void**)&s.x32, bytes32(idx, cap));
    if (e == cudaSuccess) e = cudaMalloc((void**)&s.x16, bytes16(idx, cap));
    if (e == cudaSuccess) e = cudaMalloc((void**)&s.nrm, bytesn(cap));
    if (e != cudaSuccess) {
        cudaGetLastError();
        free_set(&s);
        return 1;
    }
    s.cap = cap;
    *out = s;
    return 0;
}

static StorageSet current_set(const bfk_index* idx) {
    StorageSet s;
    s.x32 = idx->x32;
    s.x16 = idx->x16;
    s.nrm = idx->nrm;
    s.vm32 = idx->vm32;
    s.vm16 = idx->vm16;
    s.vmn = idx->vmn;
    s.cap = idx->cap;
    return s;
}

static void adopt_set(bfk_index* idx, const StorageSet& s) {
    idx->x32 = s.x32;
    idx->x16 = s.x16;
    idx->nrm = s.nrm;
    idx->vm32 = s.vm32;
    idx->vm16 = s.vm16;
    idx->vmn = s.vmn;
    idx->cap = s.cap;
}

static int free_storage(bfk_index* idx) {
    StorageSet s = current_set(idx);
    free_set(&s);
    adopt_set(idx, s);
    return 0;
}

// copies the first `rows` rows of the index into another set, on `st`
static int copy_rows_to(const bfk_index* idx, const StorageSet& dst, int64_t rows, cudaStream_t st) {
    if (rows <= 0) return 0;
    if (dst.x32) BFK_CUDA(cudaMemcpyAsync(dst.x32, idx->x32, bytes32(idx, rows), cudaMemcpyDeviceToDevice, st));
    BFK_CUDA(cudaMemcpyAsync(dst.x16, idx->x16, bytes16(idx, rows), cudaMemcpyDeviceToDevice, st));
    BFK_CUDA(cudaMemcpyAsync(dst.nrm, idx->nrm, sizeof(float) * (size_t)rows, cudaMemcpyDeviceToDevice, st));
    return 0;
}

// room for `want` rows (at least `need`): in place when the storage is growable, else a larger copy
static int grow_storage(bfk_index* idx, int64_t need, int64_t want, cudaStream_t st, const char* who) {
    if (idx->vm16 || idx->cap == 0) {
        if (idx->cap == 0 && !idx->vm16) {  // first rows: this decides between the two kinds of storage
            StorageSet s;
            if (alloc_set(idx, want, &s) && (want == need || alloc_set(idx, need, &s)))
                return fail(std::string(who) + ": out of device memory for " + std::to_string(need) + " rows");
            adopt_set(idx, s);
            return 0;
        }
        StorageSet s = current_set(idx);
        if (vm_set_grow(idx, &s, want) && (want == need || vm_set_grow(idx, &s, need)))
            return fail(std::string(who) + ": out of device memory growing the index to " + std::to_string(need) + " rows");
        idx->cap = s.cap;
        return 0;
    }
    StorageSet s;
    if (alloc_set(idx, want, &s) && (want == need || alloc_set(idx, need, &s)))
        return fail(std::string(who) + ": out of device memory growing the index to " + std::to_string(need) + " rows");
    BFK_TRY(wait_rows_ready(idx, st));
    int rc = copy_rows_to(idx, s, idx->ntotal, st);
    if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = fail(std::string(who) + ": copy failed");
    if (rc) {
        free_set(&s);
        return rc;
    }
    free_storage(idx);
    adopt_set(idx, s);
    return 0;
}

int wait_rows_ready(const bfk_index* idx, cudaStream_t st) {
    if (idx->rows_pending && idx->rows_ready && st != idx->rows_stream)
        BFK_CUDA(cudaStreamWaitEvent(st, idx->rows_ready, 0));
    return 0;
}

int mark_rows_ready(bfk_index* idx, cudaStream_t st) {
    BFK_CUDA(cudaEventRecord(idx->rows_ready, st));
    idx->rows_pending = true;
    idx->rows_stream = st;
    return 0;
}

// grow corpus storage to hold at least `need` rows; idx->mu is held.  Growable storage adds a quarter (mapping is
// cheap and nothing is copied), the cudaMalloc fall-back doubles (every regrowth copies everything)
int ensure_capacity(bfk_index* idx, int64_t need, cudaStream_t st) {
    if (need <= idx->cap) return 0;
    const int64_t want = idx->vm16 ? std::max<int64_t>(need, idx->cap + std::max<int64_t>(idx->cap / 4, 1024))
                                   : std::max<int64_t>(std::max<int64_t>(need, idx->cap * 2), 1024);
    return grow_storage(idx, need, want, st, "add");
}

int grow_for(bfk_index* idx, int64_t need, cudaStream_t st) {
    std::lock_guard<std::mutex> lk(idx->mu);
    return ensure_capacity(idx, need, st);
}

// The workspace's last user is remembered as (stream, "dirty") instead of recording an event after every search:
// the event is only recorded -- on that stream, behind the work in question -- when somebody actually has to wait for
// it (another stream uses the workspace, or it is regrown).  One driver call less on every search's critical path.
static int ws_settle(bfk_index* idx) {
    if (!idx->ws_dirty) return 0;
    idx->ws_dirty = false;
    if (cudaEventRecord(idx->ws_free, idx->ws_stream) != cudaSuccess) {
        cudaGetLastError();  // the stream is gone: everything on it has completed or will have after this
        BFK_CUDA(cudaDeviceSynchronize());
        BFK_CUDA(cudaEventRecord(idx->ws_free, 0));
    }
    return 0;
}

static int ws_acquire(bfk_index* idx, cudaStream_t st) {
    if (idx->ws_stream_known && idx->ws_stream != st) {
        BFK_TRY(ws_settle(idx));
        BFK_CUDA(cudaStreamWaitEvent(st, idx->ws_free, 0));
    }
    idx->ws_stream = st;
    idx->ws_stream_known = true;
    idx->ws_dirty = true;  // the caller is about to enqueue work on it
    return 0;
}

int ensure_workspace(bfk_index* idx, size_t bytes) {
    if (bytes <= idx->ws.bytes) return 0;
    // the previous buffer may still be in use by an enqueued search
    BFK_TRY(ws_settle(idx));
    if (idx->ws_free) BFK_CUDA(cudaEventSynchronize(idx->ws_free));
    if (idx->ws.ptr) cudaFree(idx->ws.ptr);
    idx->ws.ptr = nullptr;
    idx->ws.bytes = 0;
    size_t want = std::max(bytes, (size_t)(4u << 20));
    want = (size_t)round_up((int64_t)want, 1 << 20);
    BFK_CUDA(cudaMalloc(&idx->ws.ptr, want));
    idx->ws.bytes = want;
    return 0;
}

struct WsCarver {
    char* base;
    size_t off = 0;
    explicit WsCarver(void* p) : base((char*)p) {}
    template <typename T>
    T* take(size_t count) {
        off = (size_t)round_up((int64_t)off, 256);
        T* p = base ? (T*)(base + off) : nullptr;
        off += sizeof(T) * count;
        return p;
    }
};

static int make_index(int d, int metric, int device, int storage, bfk_index** out) {
    if (!out) return fail("null output pointer");
    *out = nullptr;
    if (d <= 0 || d > BFK_MAX_D) return fail("dimension must be in [1, " + std::to_string(BFK_MAX_D) + "]");
    if (metric != BFK_METRIC_L2 && metric != BFK_METRIC_INNER_PRODUCT)
        return fail("only METRIC_L2 and METRIC_INNER_PRODUCT flat indexes are supported");
    if (storage != BFK_STORE_F32 && storage != BFK_STORE_BF16) return fail("unknown storage mode");
    BFK_TRY(check_device(device));
    DeviceGuard g(device);
    if (!g.ok) return fail("cudaSetDevice failed");
    bfk_index* idx = new bfk_index();
    idx->d = d;
    idx->metric = metric;
    idx->device = device;
    idx->storage = storage;
    idx->ld32 = (int)round_up(d, 4);
    idx->ld16 = (int)round_up(d, 64);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) idx->num_sms = prop.multiProcessorCount;
    if (prop.major < 10) {
        delete idx;
        return fail("libbfk is built for sm_100a (B200) only; device compute capability is " +
                    std::to_string(prop.major) + "." + std::to_string(prop.minor));
    }
    if (cudaMalloc((void**)&idx->maxn, 2 * sizeof(unsigned int)) != cudaSuccess ||
        cudaMemset(idx->maxn, 0, 2 * sizeof(unsigned int)) != cudaSuccess ||
        cudaMalloc((void**)&idx->sel_ticket, sizeof(int)) != cudaSuccess ||
        cudaMemset(idx->sel_ticket, 0, sizeof(int)) != cudaSuccess ||
        cudaEventCreateWithFlags(&idx->ws_free, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&idx->rows_ready, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&idx->stage_done, cudaEventDisableTiming) != cudaSuccess ||
        cudaHostAlloc((void**)&idx->h_flag, 2 * sizeof(int), cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer((void**)&idx->h_flag_dev, idx->h_flag, 0) != cudaSuccess) {
        delete idx;
        return fail(std::string("index allocation failed: ") + cudaGetErrorString(cudaGetLastError()));
    }
    idx->h_flag[0] = idx->h_flag[1] = 0;
    *out = idx;
    return 0;
}

// ---- flat file format (faiss IxF2 / IxFI), see oracle/faiss_io.py for the layout ----
static const int64_t HEADER_BYTES = 45;

static void put_header(uint8_t* h, int d, int64_t n, int metric) {
    memcpy(h, metric == BFK_METRIC_L2 ? "IxF2" : "IxFI", 4);
    int32_t d32 = d;
    memcpy(h + 4, &d32, 4);
    memcpy(h + 8, &n, 8);
    int64_t dummy = 1 << 20;
    memcpy(h + 16, &dummy, 8);
    memcpy(h + 24, &dummy, 8);
    h[32] = 1;
    int32_t m32 = metric;
    memcpy(h + 33, &m32, 4);
    uint64_t count = (uint64_t)n * (uint64_t)d;
    memcpy(h + 37, &count, 8);
}

static int parse_header(const uint8_t* h, int64_t len, int* d, int64_t* n, int* metric, int64_t* payload_off) {
    if (len < HEADER_BYTES) return fail("flat index: file too short");
    bool f2 = !memcmp(h, "IxF2", 4), fi = !memcmp(h, "IxFI", 4), fl = !memcmp(h, "IxFl", 4);
    if (!f2 && !fi && !fl)
        return fail(std::string("Index type ") + std::string((const char*)h, 4) + " not recognized (only flat L2/IP)");
    int32_t d32, m32;
    memcpy(&d32, h + 4, 4);
    memcpy(n, h + 8, 8);
    memcpy(&m32, h + 33, 4);
    int64_t off = 37;
    if (m32 > 1) off += 4;
    if (len < off + 8) return fail("flat index: truncated header");
    uint64_t count;
    memcpy(&count, h + off, 8);
    off += 8;
    if (d32 <= 0 || *n < 0 || count != (uint64_t)(*n) * (uint64_t)d32) return fail("flat index: corrupt header");
    *d = d32;
    *metric = f2 ? BFK_METRIC_L2 : fi ? BFK_METRIC_INNER_PRODUCT : m32;
    if (*metric != BFK_METRIC_L2 && *metric != BFK_METRIC_INNER_PRODUCT)
        return fail("flat index: unsupported metric type " + std::to_string(m32));
    *payload_off = off;
    return 0;
}

// copy rows [i0, i0+n) as packed fp32 [n, d] into a host or device buffer
int launch_rows_to_f32(const bfk_index* idx, int64_t i0, int64_t n, float* out_dev, cudaStream_t st);

static int rows_out(const bfk_index* idx, int64_t i0, int64_t n, float* out, int out_is_device) {
    if (n == 0) return 0;
    cudaStream_t st = 0;
    BFK_TRY(wait_rows_ready(idx, st));  // appends are only enqueued, possibly on a non-blocking stream
    if (idx->storage == BFK_STORE_F32) {
        BFK_CUDA(cudaMemcpy2DAsync(out, sizeof(float) * idx->d, idx->x32 + (size_t)i0 * idx->ld32,
                                   sizeof(float) * idx->ld32, sizeof(float) * idx->d, (size_t)n,
                                   out_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
        BFK_CUDA(cudaStreamSynchronize(st));
        return 0;
    }
    if (out_is_device) {
        BFK_TRY(launch_rows_to_f32(idx, i0, n, out, st));
        BFK_CUDA(cudaStreamSynchronize(st));
        return 0;
    }
    const int64_t chunk = std::max<int64_t>(1, (64ll << 20) / (int64_t)(sizeof(float) * idx->d));
    float* tmp = nullptr;
    BFK_CUDA(cudaMalloc((void**)&tmp, sizeof(float) * (size_t)std::min(chunk, n) * idx->d));
    int rc = 0;
    for (int64_t j = 0; j < n && !rc; j += chunk) {
        int64_t m = std::min(chunk, n - j);
        rc = launch_rows_to_f32(idx, i0 + j, m, tmp, st);
        if (!rc && cudaMemcpyAsync(out + (size_t)j * idx->d, tmp, sizeof(float) * (size_t)m * idx->d,
                                   cudaMemcpyDeviceToHost, st) != cudaSuccess)
            rc = fail("D2H copy failed");
        if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = fail("stream sync failed");
    }
    cudaFree(tmp);
    return rc;
}

// Small host-side adds (the reference appends ONE row per call, indexer.py:41) go through the pinned staging
// buffer the searches use: a cudaMemcpyAsync from pageable memory is staged by the driver anyway, synchronously.
constexpr size_t STAGE_BYTES = 1 << 20;
static char* stage_buffer(bfk_index* idx) {
    if (!idx->h_stage && cudaHostAlloc((void**)&idx->h_stage, STAGE_BYTES, cudaHostAllocMapped) != cudaSuccess) {
        cudaGetLastError();
        idx->h_stage = nullptr;
    }
    return idx->h_stage;
}

static int add_rows(bfk_index* idx, const float* x, int64_t n, int x_is_device, cudaStream_t st) {
    if (n == 0) return 0;
    std::lock_guard<std::mutex> lk(idx->mu);  // searches of other threads see the index before or after, never during
    BFK_TRY(search_tensor_finish(idx, st, nullptr));  // a deferred search still owns the workspace
    if (!x) return fail("add: null input");
    if (idx->ntotal + n > 0x7fffffffll) return fail("add: a single shard holds at most 2^31-1 rows");
    BFK_TRY(ensure_capacity(idx, idx->ntotal + n, st));
    if (x_is_device) {
        BFK_TRY(wait_rows_ready(idx, st));  // the tail norms / maxn of an append on another stream
        BFK_TRY(launch_append_f32(idx, x, idx->d, n, st));
        idx->ntotal += n;
        return mark_rows_ready(idx, st);
    }
    // host rows: stage through device scratch in <= 64 MiB pieces
    const int64_t chunk = std::max<int64_t>(1, (64ll << 20) / (int64_t)(sizeof(float) * idx->d));
    BFK_TRY(ensure_workspace(idx, sizeof(float) * (size_t)std::min(chunk, n) * idx->d));
    BFK_TRY(ws_acquire(idx, st));
    BFK_TRY(wait_rows_ready(idx, st));
    const size_t small_bytes = sizeof(float) * (size_t)n * idx->d;
    char* pin = small_bytes <= STAGE_BYTES ? stage_buffer(idx) : nullptr;
    for (int64_t j = 0; j < n; j += chunk) {
        int64_t m = std::min(chunk, n - j);
        const void* src = x + (size_t)j * idx->d;
        if (pin) {
            memcpy(pin, src, small_bytes);
            src = pin;
        }
        BFK_CUDA(cudaMemcpyAsync(idx->ws.ptr, src, sizeof(float) * (size_t)m * idx->d, cudaMemcpyHostToDevice, st));
        if (pin) BFK_CUDA(cudaEventRecord(idx->stage_done, st));
        BFK_TRY(launch_append_f32(idx, (const float*)idx->ws.ptr, idx->d, m, st));
        idx->ntotal += m;
        // the staging buffers are reused by the next piece / the next call (and the caller owns x again on
        // return): the copy out of them must have happened.  Rows that came through the pinned buffer only wait
        // for that copy; their append kernel stays enqueued and `rows_ready` orders every later reader behind it.
        if (pin) {
            BFK_TRY(mark_rows_ready(idx, st));
            BFK_CUDA(cudaEventSynchronize(idx->stage_done));
            return 0;
        }
        BFK_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}

// ---- search orchestration ---------------------------------------------------
static int search_exact(bfk_index* idx, const float* q_dev, int ldq, int64_t nq, int k, float* D_dev,
                        int64_t* I_dev, int64_t id_offset, char* ws_tail, size_t ws_tail_bytes,
                        cudaStream_t st, bfk_search_stats* stats, cudaEvent_t ev0, cudaEvent_t ev1,
                        const SelectNotify* nt = nullptr);

}  // namespace bfk

using namespace bfk;

extern "C" {

const char* bfk_last_error(void) { return g_err.c_str(); }
int bfk_version(void) { return 100; }

int bfk_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int bfk_create(int d, int metric, int device, int storage, bfk_index** out) {
    return make_index(d, metric, device, storage, out);
}

int bfk_free(bfk_index* idx) {
    if (!idx) return 0;
    DeviceGuard g(idx->device);
    cudaDeviceSynchronize();
    search_tensor_forget(idx);
    free_storage(idx);
    if (idx->ws.ptr) cudaFree(idx->ws.ptr);
    if (idx->maxn) cudaFree(idx->maxn);
    if (idx->sel_ticket) cudaFree(idx->sel_ticket);
    if (idx->ws_free) cudaEventDestroy(idx->ws_free);
    if (idx->rows_ready) cudaEventDestroy(idx->rows_ready);
    if (idx->stage_done) cudaEventDestroy(idx->stage_done);
    if (idx->prof_ring) cudaFree(idx->prof_ring);
    if (idx->h_flag) cudaFreeHost(idx->h_flag);
    if (idx->sel_buf) cudaFree(idx->sel_buf);
    if (idx->h_stage) cudaFreeHost(idx->h_stage);
    delete idx;
    return 0;
}

int bfk_dim(const bfk_index* idx) { return idx ? idx->d : -1; }
int bfk_metric(const bfk_index* idx) { return idx ? idx->metric : -1; }
int bfk_storage(const bfk_index* idx) { return idx ? idx->storage : -1; }
int bfk_device(const bfk_index* idx) { return idx ? idx->device : -1; }
int64_t bfk_ntotal(const bfk_index* idx) { return idx ? idx->ntotal : -1; }

int bfk_reserve(bfk_index* idx, int64_t n) {
    if (!idx) return fail("null index");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    if (n <= idx->cap) return 0;
    BFK_TRY(search_tensor_finish(idx, 0, nullptr));
    // exact size: callers reserve when they know the final row count (growable storage maps up to it in place;
    // the cudaMalloc fall-back allocates it and copies what is stored, on the legacy stream)
    return grow_storage(idx, n, n, 0, "reserve");
}

int bfk_add(bfk_index* idx, const float* x, int64_t n, int x_is_device, void* stream) {
    if (!idx) return fail("null index");
    if (n < 0) return fail("add: negative row count");
    DeviceGuard g(idx->device);
    return add_rows(idx, x, n, x_is_device, (cudaStream_t)stream);
}

int bfk_add_pooled(bfk_index* idx, const void* tok, int tok_dtype, const void* mask, int mask_dtype,
                   int64_t B, int T, int H, void* stream) {
    if (!idx) return fail("null index");
    if (H != idx->d) return fail("add_pooled: hidden size != index dimension");
    if (B < 0 || T <= 0) return fail("add_pooled: bad shape");
    if (B == 0) return 0;
    if (idx->ntotal + B > 0x7fffffffll) return fail("add: a single shard holds at most 2^31-1 rows");
    DeviceGuard g(idx->device);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(idx->mu);
    BFK_TRY(search_tensor_finish(idx, st, nullptr));
    BFK_TRY(ensure_capacity(idx, idx->ntotal + B, st));
    BFK_TRY(wait_rows_ready(idx, st));
    float* o32 = idx->x32 ? idx->x32 + (size_t)idx->ntotal * idx->ld32 : nullptr;
    __nv_bfloat16* o16 = idx->x16 + (size_t)idx->ntotal * idx->ld16;
    BFK_TRY(launch_pool(tok, tok_dtype, mask, mask_dtype, B, T, H, o32, idx->ld32, o16, idx->ld16,
                        idx->nrm + idx->ntotal, idx->maxn, st));
    idx->ntotal += B;
    return mark_rows_ready(idx, st);
}

int bfk_pool_normalize(const void* tok, int tok_dtype, const void* mask, int mask_dtype, int64_t B, int T,
                       int H, float* out_f32, int io_is_device, int device, void* stream) {
    BFK_TRY(check_device(device));
    if (B < 0 || T <= 0 || H <= 0) return fail("pool_normalize: bad shape");
    if (B == 0) return 0;
    DeviceGuard g(device);
    cudaStream_t st = (cudaStream_t)stream;
    if (io_is_device)
        return launch_pool(tok, tok_dtype, mask, mask_dtype, B, T, H, out_f32, H, nullptr, 0, nullptr, nullptr, st);
    size_t tsz = tok_dtype == BFK_DT_F32 ? 4 : 2;
    size_t msz = mask_dtype == BFK_DT_I64 ? 8 : mask_dtype == BFK_DT_I32 ? 4 : 1;
    void *dtok = nullptr, *dmask = nullptr;
    float* dout = nullptr;
    int rc = 0;
    if (cudaMalloc(&dtok, tsz * (size_t)B * T * H) != cudaSuccess || cudaMalloc(&dmask, msz * (size_t)B * T) != cudaSuccess ||
        cudaMalloc((void**)&dout, sizeof(float) * (size_t)B * H) != cudaSuccess)
        rc = fail("pool_normalize: out of device memory");
    if (!rc && (cudaMemcpyAsync(dtok, tok, tsz * (size_t)B * T * H, cudaMemcpyHostToDevice, st) != cudaSuccess ||
                cudaMemcpyAsync(dmask, mask, msz * (size_t)B * T, cudaMemcpyHostToDevice, st) != cudaSuccess))
        rc = fail("pool_normalize: H2D copy failed");
    if (!rc) rc = launch_pool(dtok, tok_dtype, dmask, mask_dtype, B, T, H, dout, H, nullptr, 0, nullptr, nullptr, st);
    if (!rc && (cudaMemcpyAsync(out_f32, dout, sizeof(float) * (size_t)B * H, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                cudaStreamSynchronize(st) != cudaSuccess))
        rc = fail("pool_normalize: D2H copy failed");
    cudaFree(dtok);
    cudaFree(dmask);
    cudaFree(dout);
    return rc;
}

int bfk_search(bfk_index* idx, const float* q, int64_t nq, int64_t k, float* D, int64_t* I, int io_is_device,
               void* stream) {
    return bfk_search_ex(idx, q, nq, k, D, I, io_is_device, stream, BFK_ALGO_AUTO, 0, nullptr);
}

static int search_args_ok(bfk_index* idx, const float* q, int64_t nq, int64_t k64, float* D, int64_t* I,
                          bfk_search_stats* stats, bool* empty) {
    *empty = false;
    if (!idx) return fail("null index");
    if (nq < 0) return fail("search: negative query count");
    if (k64 <= 0) return fail("search: k must be positive");
    if (k64 > BFK_MAX_K) return fail("search: k > " + std::to_string(BFK_MAX_K) + " is not supported");
    if (stats) memset(stats, 0, sizeof(*stats));
    if (nq == 0) {
        *empty = true;
        return 0;
    }
    if (!q || !D || !I) return fail("search: null buffer");
    return 0;
}

// the search proper; idx->mu is held and the device is current.  sel_dev: optional selector bitmap on the device
static int search_locked(bfk_index* idx, const float* q, int64_t nq, int k, float* D, int64_t* I, int io_is_device,
                         cudaStream_t st, int algo, int64_t id_offset, bfk_search_stats* stats,
                         const unsigned char* sel_dev);

int bfk_search_ex(bfk_index* idx, const float* q, int64_t nq, int64_t k64, float* D, int64_t* I,
                  int io_is_device, void* stream, int algo, int64_t id_offset, bfk_search_stats* stats) {
    bool empty;
    BFK_TRY(search_args_ok(idx, q, nq, k64, D, I, stats, &empty));
    if (empty) return 0;
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    return search_locked(idx, q, nq, (int)k64, D, I, io_is_device, (cudaStream_t)stream, algo, id_offset, stats, nullptr);
}

int bfk_search_sel(bfk_index* idx, const float* q, int64_t nq, int64_t k64, float* D, int64_t* I, int io_is_device,
                   void* stream, int algo, int64_t id_offset, const uint8_t* bitmap, int bitmap_is_device,
                   bfk_search_stats* stats) {
    bool empty;
    BFK_TRY(search_args_ok(idx, q, nq, k64, D, I, stats, &empty));
    if (empty) return 0;
    if (!bitmap) return fail("search_sel: null selector bitmap");
    DeviceGuard g(idx->device);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(idx->mu);
    BFK_TRY(search_tensor_finish(idx, st, nullptr));
    // selector staging: [bitmap, ceil(n/8) bytes rounded to 256][bias, tiles*256 floats]
    const size_t bm_bytes = (size_t)((idx->ntotal + 7) / 8);
    const size_t bm_pad = (size_t)round_up((int64_t)bm_bytes + 1, 256);
    const int64_t n_pad = round_up(std::max<int64_t>(idx->ntotal, 1), 256);
    const size_t need = bm_pad + sizeof(float) * (size_t)n_pad;
    if (idx->sel_buf_bytes < need) {
        BFK_CUDA(cudaStreamSynchronize(st));
        if (idx->sel_buf) cudaFree(idx->sel_buf);
        idx->sel_buf = nullptr;
        idx->sel_buf_bytes = 0;
        BFK_CUDA(cudaMalloc(&idx->sel_buf, need));
        idx->sel_buf_bytes = need;
    }
    unsigned char* bm_dev = (unsigned char*)idx->sel_buf;
    BFK_CUDA(cudaMemcpyAsync(bm_dev, bitmap, bm_bytes, bitmap_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    int rc = search_locked(idx, q, nq, (int)k64, D, I, io_is_device, st, algo & ~BFK_ALGO_FLAG_DEFER, id_offset, stats, bm_dev);
    idx->sel = nullptr;
    idx->sel_bias = nullptr;
    return rc;
}

static int search_locked(bfk_index* idx, const float* q, int64_t nq, int k, float* D, int64_t* I, int io_is_device,
                         cudaStream_t st, int algo, int64_t id_offset, bfk_search_stats* stats,
                         const unsigned char* sel_dev) {
    const int defer = (algo & BFK_ALGO_FLAG_DEFER) && io_is_device && !stats && !sel_dev;
    algo &= ~BFK_ALGO_FLAG_DEFER;
    BFK_TRY(search_tensor_finish(idx, st, nullptr));  // a deferred search nobody settled: settle it first
    BFK_TRY(wait_rows_ready(idx, st));                 // rows appended on another stream are only enqueued there

    if (algo == BFK_ALGO_AUTO) {
        // The tensor path wins at every batch size, one query included: it streams the bf16 scan copy (half the
        // bytes of the fp32 rows: nq=1 0.168 ms against 0.248 ms, nq=8 0.171 against 0.457 on 1M x 384) and its
        // answers are re-scored in fp32 and certified exact.  The fp32 scan remains the explicit
        // BFK_ALGO_EXACT_SIMT, the fallback for uncertified queries, and the path for shapes the tensor path
        // does not take (k > 128).
        algo = tensor_path_supported(idx, nq, k) ? BFK_ALGO_TENSOR : BFK_ALGO_EXACT_SIMT;
        // ...except for a handful of queries over a small fp32 corpus, where the tensor path's fixed cost (one more
        // launch, tensor maps, the certificate read-back: ~45 us against ~10) outweighs the bytes it saves.  The
        // reference's own index (649 x 384) answers in 53 us through the fp32 scan, 97 us through the tensor path;
        // the two meet near 250 k rows at d = 384 for one or two queries, earlier once the fp32 scan turns FMA-bound.
        if (algo == BFK_ALGO_TENSOR && idx->storage == BFK_STORE_F32 && nq <= 8 &&
            idx->ntotal * (int64_t)idx->d <= (nq <= 2 ? (int64_t)64 << 20 : (int64_t)24 << 20))
            algo = BFK_ALGO_EXACT_SIMT;
    } else if (algo == BFK_ALGO_TENSOR) {
        if (!tensor_path_supported(idx, nq, k)) return fail("search: tensor path does not support this shape");
    } else if (algo != BFK_ALGO_EXACT_SIMT) {
        return fail("search: unknown algorithm");
    }

    // workspace: [queries fp32 padded][D][I][algorithm scratch...]
    const int ldq = idx->ld32;
    size_t scratch = 0;
    int nparts = 0;
    BFK_TRY(scan_exact_workspace(idx, (int)std::min<int64_t>(nq, 8), k, &scratch, &nparts));
    size_t exact_bytes = sizeof(uint64_t) * (size_t)nq * nparts * k;
    size_t tail_bytes = exact_bytes;
    if (algo == BFK_ALGO_TENSOR) {
        // the tensor path sizes its own scratch inside search_tensor via ensure_workspace's tail
        tail_bytes = std::max(tail_bytes, tensor_workspace_bytes(idx, nq, k));
    }
    WsCarver sizing(nullptr);
    sizing.take<float>((size_t)nq * ldq);
    sizing.take<float>((size_t)nq * k);
    sizing.take<int64_t>((size_t)nq * k);
    sizing.take<char>(256);
    size_t head = sizing.off;
    BFK_TRY(ensure_workspace(idx, head + tail_bytes + 512));
    // the workspace's previous user: work on the same stream is ordered anyway, another stream has to be waited for
    BFK_TRY(ws_acquire(idx, st));
    WsCarver ws(idx->ws.ptr);
    float* q_dev = ws.take<float>((size_t)nq * ldq);
    float* D_dev = ws.take<float>((size_t)nq * k);
    int64_t* I_dev = ws.take<int64_t>((size_t)nq * k);
    char* tail = ws.take<char>(256);
    size_t tail_avail = idx->ws.bytes - (size_t)(tail - (char*)idx->ws.ptr);

    // How host-side queries get in and results get out.
    //   PINNED  the caller's buffer is page-locked (cudaHostAlloc / cudaHostRegister / torch pin_memory): the kernels
    //           address it directly -- the query preparation reads it over PCIe, K5 writes (D, I) into it;
    //   STAGED  pageable and small: one memcpy into / out of the index's own pinned buffer, which the kernels address
    //           the same way (queries in its first half, results in its second);
    //   COPY    pageable and large (or a timed / selector search): cudaMemcpyAsync on the stream, one stream wait at the end.
    // With PINNED / STAGED nothing is enqueued on the stream but the kernels, and the host only polls the word the last
    // CTA of K5 (tensor path) or of K3a / K4 (fp32 path) writes: two DMA launches, their serialisation with the kernels
    // and a driver-level wait less.
    enum Mode { MODE_DEVICE, MODE_PINNED, MODE_STAGED, MODE_COPY };
    const size_t q_bytes = sizeof(float) * (size_t)nq * idx->d;
    const size_t out_span = (size_t)((char*)(I_dev + (size_t)nq * k) - (char*)D_dev);  // D, padding, I: contiguous
    const bool direct_ok = !stats && !sel_dev;  // timed and selector searches go through copies
    Mode q_mode = MODE_DEVICE, out_mode = MODE_DEVICE;
    const float* q_direct = nullptr;  // queries as the GPU addresses them when q_mode is PINNED / STAGED
    char* stage_out = nullptr;
    if (!io_is_device) {
        q_mode = out_mode = MODE_COPY;
        auto pinned = [](const void* p, void** dev) {
            cudaPointerAttributes a;
            if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
                cudaGetLastError();
                return false;
            }
            *dev = a.devicePointer;
            return a.type == cudaMemoryTypeHost && a.devicePointer != nullptr;
        };
        void *qd = nullptr, *Dd = nullptr, *Id = nullptr;
        // a few KB are cheaper to memcpy through the staging buffer than to ask the driver about (~1 us per pointer):
        // the reference's own call is one 384-float query and 3 results (llm-qa/main.py:101)
        const bool tiny = q_bytes <= 16384 && out_span <= 16384 && stage_buffer(idx) != nullptr;
        if (tiny) {
            q_mode = out_mode = MODE_STAGED;
            q_direct = reinterpret_cast<const float*>(idx->h_stage);
            stage_out = idx->h_stage + STAGE_BYTES / 2;
        } else if (direct_ok && pinned(q, &qd)) {
            q_mode = MODE_PINNED;
            q_direct = static_cast<const float*>(qd);
        } else if (q_bytes <= STAGE_BYTES / 2 && stage_buffer(idx)) {
            q_mode = MODE_STAGED;
            q_direct = reinterpret_cast<const float*>(idx->h_stage);
        }
        if (tiny) {
        } else if (direct_ok && pinned(D, &Dd) && pinned(I, &Id)) out_mode = MODE_PINNED;
        else if (out_span <= STAGE_BYTES / 2 && stage_buffer(idx)) {
            out_mode = MODE_STAGED;
            stage_out = idx->h_stage + STAGE_BYTES / 2;
        }
        if (out_mode == MODE_PINNED) {
            D = static_cast<float*>(Dd);
            I = static_cast<int64_t*>(Id);
        }
    }
    // the fp32 scan reads the queries once per CTA: over PCIe that only pays while (queries x CTAs) stays small --
    // the reference's 649-row index is 21 CTAs x 1.5 KB; larger scans get their queries by one DMA copy instead
    const bool q_direct_ok = direct_ok && (algo == BFK_ALGO_TENSOR || q_bytes * (size_t)std::max(nparts, 1) <= 65536);
    const bool q_zero_copy = q_direct_ok && (q_mode == MODE_PINNED || q_mode == MODE_STAGED);
    const bool out_zero_copy = direct_ok && (out_mode == MODE_PINNED || out_mode == MODE_STAGED);
    if (q_mode == MODE_STAGED) memcpy(idx->h_stage, q, q_bytes);
    if (q_zero_copy) {
        // prep_queries_kernel reads q_direct and leaves a padded fp32 copy in q_dev for K5
    } else if (q_mode == MODE_STAGED) {
        if (ldq != idx->d) BFK_CUDA(cudaMemsetAsync(q_dev, 0, sizeof(float) * (size_t)nq * ldq, st));
        BFK_CUDA(cudaMemcpy2DAsync(q_dev, sizeof(float) * ldq, idx->h_stage, sizeof(float) * idx->d, sizeof(float) * idx->d,
                                   (size_t)nq, cudaMemcpyHostToDevice, st));
    } else if (q_mode == MODE_PINNED) {
        if (ldq != idx->d) BFK_CUDA(cudaMemsetAsync(q_dev, 0, sizeof(float) * (size_t)nq * ldq, st));
        BFK_CUDA(cudaMemcpy2DAsync(q_dev, sizeof(float) * ldq, q, sizeof(float) * idx->d, sizeof(float) * idx->d, (size_t)nq,
                                   cudaMemcpyHostToDevice, st));
    } else if (io_is_device && ldq == idx->d && !defer && (reinterpret_cast<uintptr_t>(q) & 15u) == 0) {
        // device queries whose row stride already is the padded one: read them in place (a deferred search
        // keeps its own copy, its retry may run after the caller has let go of the buffer)
        q_dev = const_cast<float*>(q);
    } else {
        if (ldq != idx->d) BFK_CUDA(cudaMemsetAsync(q_dev, 0, sizeof(float) * (size_t)nq * ldq, st));
        BFK_CUDA(cudaMemcpy2DAsync(q_dev, sizeof(float) * ldq, q, sizeof(float) * idx->d, sizeof(float) * idx->d,
                                   (size_t)nq, io_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    }
    float* D_out = io_is_device ? D : D_dev;
    int64_t* I_out = io_is_device ? I : I_dev;
    if (out_zero_copy && out_mode == MODE_PINNED) {
        D_out = D;
        I_out = I;
    } else if (out_zero_copy) {  // same layout as the device pair: D, padding, I
        D_out = reinterpret_cast<float*>(stage_out);
        I_out = reinterpret_cast<int64_t*>(stage_out + ((char*)I_dev - (char*)D_dev));
    }

    if (sel_dev) {
        idx->sel = sel_dev;
        if (algo == BFK_ALGO_TENSOR) {
            const size_t bm_pad = (size_t)round_up((idx->ntotal + 7) / 8 + 1, 256);
            float* bias = reinterpret_cast<float*>((char*)idx->sel_buf + bm_pad);
            const int64_t n_pad = round_up(std::max<int64_t>(idx->ntotal, 1), 256);
            BFK_TRY(launch_sel_bias(sel_dev, idx->nrm, idx->ntotal, n_pad, idx->metric == BFK_METRIC_L2 ? 1 : 0, bias, st));
            idx->sel_bias = bias;
        }
    }

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    if (stats) {
        BFK_CUDA(cudaEventCreate(&ev0));
        BFK_CUDA(cudaEventCreate(&ev1));
    }
    // A host-side tensor search is enqueued like a deferred one: the (D, I) copy-out queues up behind K5 and the
    // host waits ONCE, then looks at the certificate count (a re-run of uncertified queries repeats the copy-out)
    // -- instead of waiting for the count first and for the copy afterwards.
    const bool host_defer = !io_is_device && algo == BFK_ALGO_TENSOR && !stats && !sel_dev;
    int rc;
    if (algo == BFK_ALGO_TENSOR) {
        rc = search_tensor(idx, q_dev, ldq, nq, k, D_out, I_out, id_offset, tail, tail_avail, st, stats, ev0, ev1,
                           (defer || host_defer) ? 1 : 0, q_zero_copy ? q_direct : nullptr, out_zero_copy ? 1 : 0);
    } else {
        // the fp32 scan reads the queries where they are (pinned host memory included: one read per CTA at kernel
        // start, row stride d), K4 writes (D, I) to the host and publishes a sequence number the host polls
        SelectNotify nt{idx->sel_ticket, idx->h_flag_dev + 1, 0};
        if (out_zero_copy) nt.seq = ++idx->flag_seq;
        rc = search_exact(idx, q_zero_copy ? q_direct : q_dev, q_zero_copy ? idx->d : ldq, nq, k, D_out, I_out, id_offset,
                          tail, tail_avail, st, stats, ev0, ev1, out_zero_copy ? &nt : nullptr);
    }
    const bool staged_out = !io_is_device && out_mode == MODE_STAGED;
    auto copy_out = [&]() -> int {
        if (out_zero_copy) return 0;  // K5 (or the fall-back's copies) wrote the host buffer directly
        if (staged_out) {
            if (cudaMemcpyAsync(stage_out, D_dev, out_span, cudaMemcpyDeviceToHost, st) != cudaSuccess)
                return fail("search: D2H copy failed");
        } else if (cudaMemcpyAsync(D, D_dev, sizeof(float) * (size_t)nq * k, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                   cudaMemcpyAsync(I, I_dev, sizeof(int64_t) * (size_t)nq * k, cudaMemcpyDeviceToHost, st) != cudaSuccess) {
            return fail("search: D2H copy failed");
        }
        return 0;
    };
    if (!rc && !io_is_device) rc = copy_out();
    if (!rc && (!io_is_device || stats) && !out_zero_copy) {
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) rc = fail(std::string("search: kernel execution failed: ") + cudaGetErrorString(e));
        if (!rc && st == idx->rows_stream) idx->rows_pending = false;
    }
    if (!rc && host_defer) {
        int64_t rewritten = 0;
        // reads the count (zero copy: polls K5's word in pinned memory instead of waiting for the stream), re-runs if needed
        rc = search_tensor_finish(idx, st, &rewritten);
        if (!rc && rewritten > 0) {
            rc = copy_out();
            if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = fail("search: kernel execution failed");
        }
    }
    if (!rc && out_zero_copy && algo == BFK_ALGO_EXACT_SIMT) rc = wait_flag_seq(idx, st);
    if (!rc && staged_out) {
        memcpy(D, stage_out, sizeof(float) * (size_t)nq * k);
        memcpy(I, stage_out + ((char*)I_dev - (char*)D_dev), sizeof(int64_t) * (size_t)nq * k);
    }
    if (stats) {
        if (!rc) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, ev0, ev1) == cudaSuccess) stats->scan_ms = ms;
            stats->algo_used = algo;
        }
        cudaEventDestroy(ev0);
        cudaEventDestroy(ev1);
    }
    return rc;
}

int bfk_search_finish(bfk_index* idx, void* stream, int64_t* n_rewritten) {
    if (!idx) return fail("null index");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    return search_tensor_finish(idx, (cudaStream_t)stream, n_rewritten);
}

int bfk_profile_begin(bfk_index* idx) {
    if (!idx) return fail("null index");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    constexpr int64_t CAP = 1 << 16;
    if (!idx->prof_ring) {
        BFK_CUDA(cudaMalloc((void**)&idx->prof_ring, sizeof(unsigned long long) * 4 * CAP));
        idx->prof_cap = CAP;
    }
    // per pass: [0] first scan CTA in (atomicMin), [1] last scan CTA out (atomicMax),
    //           [2] prep kernel in (atomicMin),    [3] last K5 CTA out (atomicMax)
    std::vector<unsigned long long> init((size_t)4 * idx->prof_cap);
    for (int64_t i = 0; i < idx->prof_cap; i++) {
        init[(size_t)4 * i] = init[(size_t)4 * i + 2] = ~0ull;
        init[(size_t)4 * i + 1] = init[(size_t)4 * i + 3] = 0;
    }
    BFK_CUDA(cudaMemcpy(idx->prof_ring, init.data(), init.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    idx->prof_used = 0;
    idx->prof_on = true;
    return 0;
}

int bfk_profile_end(bfk_index* idx, double* scan_ms_sum, double* pass_ms_sum, double* idle_ms_sum, int64_t* n_scans) {
    if (!idx) return fail("null index");
    if (scan_ms_sum) *scan_ms_sum = 0.0;
    if (pass_ms_sum) *pass_ms_sum = 0.0;
    if (idle_ms_sum) *idle_ms_sum = 0.0;
    if (n_scans) *n_scans = 0;
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->prof_on) return fail("profile_end without profile_begin");
    idx->prof_on = false;
    BFK_CUDA(cudaDeviceSynchronize());
    std::vector<unsigned long long> h((size_t)4 * idx->prof_used);
    if (idx->prof_used)
        BFK_CUDA(cudaMemcpy(h.data(), idx->prof_ring, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    double scan_ns = 0.0, pass_ns = 0.0, idle_ns = 0.0;
    int64_t n = 0;
    unsigned long long prev_end = 0;
    for (int64_t i = 0; i < idx->prof_used; i++) {
        const unsigned long long* t = &h[(size_t)4 * i];
        if (t[0] == ~0ull || t[1] < t[0]) continue;
        scan_ns += (double)(t[1] - t[0]);
        n++;
        if (t[2] != ~0ull && t[3] >= t[2]) {
            pass_ns += (double)(t[3] - t[2]);
            if (prev_end && t[2] >= prev_end) idle_ns += (double)(t[2] - prev_end);
            prev_end = t[3];
        }
    }
    if (scan_ms_sum) *scan_ms_sum = scan_ns * 1e-6;
    if (pass_ms_sum) *pass_ms_sum = pass_ns * 1e-6;
    if (idle_ms_sum) *idle_ms_sum = idle_ns * 1e-6;
    if (n_scans) *n_scans = n;
    return 0;
}

int64_t bfk_profile_stamps(bfk_index* idx, unsigned long long* out, int64_t max_passes) {
    if (!idx || !out || max_passes < 0) return -1;
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    const int64_t n = std::min(max_passes, idx->prof_used);
    if (n > 0 && (cudaDeviceSynchronize() != cudaSuccess ||
                  cudaMemcpy(out, idx->prof_ring, sizeof(unsigned long long) * 4 * (size_t)n, cudaMemcpyDeviceToHost) != cudaSuccess)) {
        cudaGetLastError();
        return -1;
    }
    return n;
}

const void* bfk_search_pending_nflag_ptr(bfk_index* idx) { return idx ? (const void*)search_tensor_pending_nflag(idx) : nullptr; }

int bfk_reset(bfk_index* idx) {
    if (!idx) return fail("null index");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    BFK_TRY(search_tensor_finish(idx, 0, nullptr));
    BFK_CUDA(cudaDeviceSynchronize());
    idx->rows_pending = false;
    idx->ntotal = 0;
    BFK_CUDA(cudaMemset(idx->maxn, 0, 2 * sizeof(unsigned int)));
    return 0;
}

int bfk_reconstruct(const bfk_index* idx, int64_t i0, int64_t n, float* out, int out_is_device) {
    if (!idx) return fail("null index");
    if (n < 0 || i0 < 0 || i0 + n > idx->ntotal) return fail("reconstruct: id range out of bounds");
    if (n == 0) return 0;
    if (!out) return fail("reconstruct: null output");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    return rows_out(idx, i0, n, out, out_is_device);
}

int bfk_remove_ids(bfk_index* idx, const int64_t* ids, int64_t n, int64_t* nremoved) {
    if (!idx) return fail("null index");
    if (nremoved) *nremoved = 0;
    if (n < 0 || (n > 0 && !ids)) return fail("remove_ids: bad arguments");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    std::vector<int64_t> rm;
    rm.reserve((size_t)n);
    for (int64_t i = 0; i < n; i++)
        if (ids[i] >= 0 && ids[i] < idx->ntotal) rm.push_back(ids[i]);
    std::sort(rm.begin(), rm.end());
    rm.erase(std::unique(rm.begin(), rm.end()), rm.end());
    if (rm.empty()) return 0;
    BFK_TRY(search_tensor_finish(idx, 0, nullptr));
    BFK_CUDA(cudaDeviceSynchronize());
    idx->rows_pending = false;
    // IndexFlatCodes::remove_ids keeps the surviving rows in order: move each kept segment down
    // through a fresh set of arrays (device copies must not overlap)
    const int64_t left = idx->ntotal - (int64_t)rm.size();
    const int64_t newcap = std::max<int64_t>(left, 1024);
    StorageSet ns;
    if (alloc_set(idx, newcap, &ns)) return fail("remove_ids: out of device memory");
    int64_t dst = 0, src = 0;
    cudaStream_t st = 0;
    auto copy_seg = [&](int64_t s, int64_t e) -> int {
        int64_t m = e - s;
        if (m <= 0) return 0;
        if (ns.x32) BFK_CUDA(cudaMemcpyAsync(ns.x32 + (size_t)dst * idx->ld32, idx->x32 + (size_t)s * idx->ld32,
                                             sizeof(float) * (size_t)m * idx->ld32, cudaMemcpyDeviceToDevice, st));
        BFK_CUDA(cudaMemcpyAsync(ns.x16 + (size_t)dst * idx->ld16, idx->x16 + (size_t)s * idx->ld16,
                                 sizeof(__nv_bfloat16) * (size_t)m * idx->ld16, cudaMemcpyDeviceToDevice, st));
        BFK_CUDA(cudaMemcpyAsync(ns.nrm + dst, idx->nrm + s, sizeof(float) * (size_t)m, cudaMemcpyDeviceToDevice, st));
        dst += m;
        return 0;
    };
    int rc = 0;
    for (size_t r = 0; r < rm.size() && !rc; r++) {
        rc = copy_seg(src, rm[r]);
        src = rm[r] + 1;
    }
    if (!rc) rc = copy_seg(src, idx->ntotal);
    if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = fail("remove_ids: copy failed");
    if (rc) {
        free_set(&ns);
        return rc;
    }
    free_storage(idx);
    adopt_set(idx, ns);
    idx->ntotal = left;
    // the tensor maps describe the old arrays
    idx->tmap[0].base = nullptr;
    idx->tmap[1].base = nullptr;
    if (nremoved) *nremoved = (int64_t)rm.size();
    return 0;
}

int64_t bfk_serialized_size(const bfk_index* idx) {
    if (!idx) return -1;
    return HEADER_BYTES + (int64_t)sizeof(float) * idx->ntotal * idx->d;
}

int bfk_serialize(const bfk_index* idx, uint8_t* buf, int64_t buf_len) {
    if (!idx || !buf) return fail("serialize: null argument");
    if (buf_len < bfk_serialized_size(idx)) return fail("serialize: buffer too small");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    put_header(buf, idx->d, idx->ntotal, idx->metric);
    return rows_out(idx, 0, idx->ntotal, (float*)(buf + HEADER_BYTES), 0);
}

int bfk_deserialize(const uint8_t* buf, int64_t len, int device, int storage, bfk_index** out) {
    if (!buf || !out) return fail("deserialize: null argument");
    int d, metric;
    int64_t n, off;
    BFK_TRY(parse_header(buf, len, &d, &n, &metric, &off));
    if (len < off + (int64_t)sizeof(float) * n * d) return fail("flat index: truncated payload");
    BFK_TRY(make_index(d, metric, device, storage, out));
    DeviceGuard g(device);
    // payload starts at byte 45: not 4-byte aligned in general, so bounce through an aligned copy
    std::vector<float> tmp((size_t)n * d);
    memcpy(tmp.data(), buf + off, sizeof(float) * (size_t)n * d);
    int rc = add_rows(*out, tmp.data(), n, 0, 0);
    if (rc) {
        bfk_free(*out);
        *out = nullptr;
    }
    return rc;
}

int bfk_write(const bfk_index* idx, const char* path) {
    if (!idx || !path) return fail("write_index: null argument");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    FILE* f = fopen(path, "wb");
    if (!f) return fail(std::string("could not open ") + path + " for writing");
    uint8_t h[HEADER_BYTES];
    put_header(h, idx->d, idx->ntotal, idx->metric);
    int rc = fwrite(h, 1, HEADER_BYTES, f) == (size_t)HEADER_BYTES ? 0 : fail("write_index: short write");
    const int64_t chunk = std::max<int64_t>(1, (256ll << 20) / (int64_t)(sizeof(float) * idx->d));
    std::vector<float> tmp((size_t)std::min(chunk, std::max<int64_t>(idx->ntotal, 1)) * idx->d);
    for (int64_t j = 0; j < idx->ntotal && !rc; j += chunk) {
        int64_t m = std::min(chunk, idx->ntotal - j);
        rc = rows_out(idx, j, m, tmp.data(), 0);
        if (!rc && fwrite(tmp.data(), sizeof(float), (size_t)m * idx->d, f) != (size_t)m * idx->d)
            rc = fail("write_index: short write");
    }
    if (fclose(f) != 0 && !rc) rc = fail("write_index: close failed");
    return rc;
}

// indexer.py:26-30 rewrites the whole file after every message; this appends the rows the file does not hold yet.
// Order of writes: payload of the new rows past the end of the old payload (a reader of the old header never
// looks there), flush, then the 45-byte header with the new ntotal / count in ONE write -- the commit point.
int bfk_write_append(bfk_index* idx, const char* path, int64_t rows_on_disk) {
    if (!idx || !path) return fail("write_index_append: null argument");
    DeviceGuard g(idx->device);
    std::lock_guard<std::mutex> lk(idx->mu);
    if (rows_on_disk < 0 || rows_on_disk > idx->ntotal)
        return fail("write_index_append: the file is said to hold " + std::to_string(rows_on_disk) + " rows, the index has " +
                    std::to_string(idx->ntotal));
    FILE* f = fopen(path, "r+b");
    if (!f) return fail(std::string("could not open ") + path + " for appending");
    uint8_t h[HEADER_BYTES + 4];
    const size_t got = fread(h, 1, sizeof(h), f);
    int d, metric;
    int64_t n, off;
    int rc = parse_header(h, (int64_t)got, &d, &n, &metric, &off);
    if (!rc && (d != idx->d || metric != idx->metric || n != rows_on_disk || off != HEADER_BYTES))
        rc = fail("write_index_append: the file's header (d=" + std::to_string(d) + ", ntotal=" + std::to_string(n) +
                  ") does not describe the first " + std::to_string(rows_on_disk) + " rows of this index");
    const int64_t m_total = idx->ntotal - rows_on_disk;
    if (!rc && m_total > 0) {
        if (fseeko(f, (off_t)(HEADER_BYTES + (int64_t)sizeof(float) * rows_on_disk * idx->d), SEEK_SET) != 0)
            rc = fail("write_index_append: seek failed");
        // the tail leaves the GPU through the pinned staging buffer when it fits (a message is a few rows)
        const size_t row_bytes = sizeof(float) * (size_t)idx->d;
        char* pin = (size_t)m_total * row_bytes <= STAGE_BYTES ? stage_buffer(idx) : nullptr;
        const int64_t chunk = pin ? m_total : std::max<int64_t>(1, (int64_t)((256ull << 20) / row_bytes));
        std::vector<float> tmp;
        if (!pin) tmp.resize((size_t)std::min(chunk, m_total) * idx->d);
        float* buf = pin ? reinterpret_cast<float*>(pin) : tmp.data();
        for (int64_t j = 0; j < m_total && !rc; j += chunk) {
            const int64_t m = std::min(chunk, m_total - j);
            rc = rows_out(idx, rows_on_disk + j, m, buf, 0);
            if (!rc && fwrite(buf, sizeof(float), (size_t)m * idx->d, f) != (size_t)m * idx->d)
                rc = fail("write_index_append: short write");
        }
        // debris of an append that never committed may reach further than this one: the file ends with its rows
        if (!rc && (fflush(f) != 0 ||
                    ftruncate(fileno(f), (off_t)(HEADER_BYTES + (int64_t)sizeof(float) * idx->ntotal * idx->d)) != 0 ||
                    fdatasync(fileno(f)) != 0))
            rc = fail("write_index_append: flush failed");
        if (!rc) {
            uint8_t nh[HEADER_BYTES];
            put_header(nh, idx->d, idx->ntotal, idx->metric);
            if (pwrite(fileno(f), nh, HEADER_BYTES, 0) != (ssize_t)HEADER_BYTES) rc = fail("write_index_append: header write failed");
            else if (fdatasync(fileno(f)) != 0) rc = fail("write_index_append: flush failed");
        }
    }
    if (fclose(f) != 0 && !rc) rc = fail("write_index_append: close failed");
    return rc;
}

int bfk_read(const char* path, int device, int storage, bfk_index** out) {
    if (!path || !out) return fail("read_index: null argument");
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(std::string("could not open ") + path + " for reading");
    uint8_t h[HEADER_BYTES + 4];
    size_t got = fread(h, 1, sizeof(h), f);
    int d, metric;
    int64_t n, off;
    int rc = parse_header(h, (int64_t)got, &d, &n, &metric, &off);
    if (!rc) rc = make_index(d, metric, device, storage, out);
    if (!rc) {
        DeviceGuard g(device);
        fseek(f, (long)off, SEEK_SET);
        const int64_t chunk = std::max<int64_t>(1, (256ll << 20) / (int64_t)(sizeof(float) * d));
        std::vector<float> tmp((size_t)std::min(chunk, std::max<int64_t>(n, 1)) * d);
        rc = bfk_reserve(*out, std::max<int64_t>(n, 1));
        for (int64_t j = 0; j < n && !rc; j += chunk) {
            int64_t m = std::min(chunk, n - j);
            if (fread(tmp.data(), sizeof(float), (size_t)m * d, f) != (size_t)m * d)
                rc = fail("read_index: truncated payload");
            if (!rc) rc = add_rows(*out, tmp.data(), m, 0, 0);
        }
        if (rc) {
            bfk_free(*out);
            *out = nullptr;
        }
    }
    fclose(f);
    return rc;
}

}  // extern "C"

namespace bfk {

// Exact path: groups of <= 8 queries, one corpus pass per group, then one select launch.
static int search_exact(bfk_index* idx, const float* q_dev, int ldq, int64_t nq, int k, float* D_dev,
                        int64_t* I_dev, int64_t id_offset, char* ws_tail, size_t ws_tail_bytes,
                        cudaStream_t st, bfk_search_stats* stats, cudaEvent_t ev0, cudaEvent_t ev1,
                        const SelectNotify* nt) {
    size_t scratch = 0;
    int nparts = 0;
    BFK_TRY(scan_exact_workspace(idx, (int)std::min<int64_t>(nq, 8), k, &scratch, &nparts));
    size_t need = sizeof(uint64_t) * (size_t)nq * nparts * k;
    if (need > ws_tail_bytes) return fail("internal: exact-scan workspace under-sized");
    uint64_t* part = (uint64_t*)ws_tail;
    int launches = 0;
    if (stats) cudaEventRecord(ev0, st);
    bool fused = false;
    if (nq <= 8 && !stats) {
        // one corpus pass: its last CTA may merge the per-CTA lists itself (decided by launch_scan_exact)
        ScanFuse sf{D_dev, I_dev, (int64_t)k, id_offset, idx->sel_ticket, nt ? nt->h_seq : nullptr, nt ? nt->seq : 0};
        BFK_TRY(launch_scan_exact(idx, q_dev, ldq, (int)nq, k, part, nparts, st, idx->sel, &sf, &fused));
        launches++;
    } else {
        for (int64_t q0 = 0; q0 < nq;) {
            int g = (int)std::min<int64_t>(nq - q0, 8);
            BFK_TRY(launch_scan_exact(idx, q_dev + (size_t)q0 * ldq, ldq, g, k, part + (size_t)q0 * nparts * k, nparts, st, idx->sel));
            launches++;
            q0 += g;
        }
    }
    if (stats) cudaEventRecord(ev1, st);
    if (!fused) {
        BFK_TRY(launch_select(part, (int64_t)nparts * k, nq, k, idx->metric, id_offset, D_dev, I_dev, k, st, nt));
        launches++;
    }
    if (stats) {
        stats->n_launches = launches;
        stats->algo_used = BFK_ALGO_EXACT_SIMT;
    }
    return 0;
}

}  // namespace bfk

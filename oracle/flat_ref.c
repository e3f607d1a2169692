/*
 * ORACLE (test infrastructure, not product code): C restatement of the CPU
 * flat kNN that the reference runs inside faiss-cpu.
 *
 * PARITY UNPINNED against live faiss (third-party, unpinned in
 * semantic-indexer/requirements.txt:2, absent from /root/reference and from
 * this image).  Restates the published algorithm behind the reference call
 * sites  indexer.py:41 (index.add)  and  llm-qa/main.py:101,117
 * (index.search via LangChain):
 *
 *   - flatref_knn_seq      : IndexFlat::search for nq < 20
 *                            (exhaustive_L2sqr_seq / exhaustive_inner_product_seq:
 *                            one OpenMP iteration per query, fvec_L2sqr /
 *                            fvec_inner_product per row, k-heap, heap_reorder)
 *   - flatref_block_update : the HeapBlockResultHandler step of the nq >= 20
 *                            BLAS path (the sgemm itself is done by numpy's
 *                            OpenBLAS in oracle/cpu_baseline.py)
 *   - flatref_pool_normalize: sentence-transformers Pooling(mean)+Normalize
 *                            (indexer.py:37), formulas in oracle/pool_ref.py
 *
 * It is validated against oracle/flat_ref.py (numpy) in tests/ and is the
 * timed "port" CPU baseline of bench.py.  Nothing in the product path links it.
 *
 * Build: oracle/Makefile  ->  oracle/libflat_ref.so
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define METRIC_IP 0
#define METRIC_L2 1

/* ---- distances: fvec_L2sqr / fvec_inner_product ---- */
static inline float l2sqr(const float* x, const float* y, int d) {
    float s = 0.f;
#pragma omp simd reduction(+ : s)
    for (int i = 0; i < d; i++) {
        float t = x[i] - y[i];
        s += t * t;
    }
    return s;
}

static inline float inner(const float* x, const float* y, int d) {
    float s = 0.f;
#pragma omp simd reduction(+ : s)
    for (int i = 0; i < d; i++) s += x[i] * y[i];
    return s;
}

/* ---- k-heap on keys (smaller key = better); root = worst kept ----
 * key = distance for L2, -inner product for IP.  Ordering is on (key, id) so
 * that equal keys resolve by id, as faiss >= 1.7.3 does (CMax::cmp2). */
static inline int worse(float ka, int64_t ia, float kb, int64_t ib) {
    return (ka > kb) || (ka == kb && ia > ib);
}

static void heap_replace_top(int64_t k, float* hk, int64_t* hi, float key, int64_t id) {
    int64_t i = 0;
    for (;;) {
        int64_t l = 2 * i + 1, r = l + 1, c;
        if (l >= k) break;
        c = (r < k && worse(hk[r], hi[r], hk[l], hi[l])) ? r : l;
        if (!worse(hk[c], hi[c], key, id)) break;
        hk[i] = hk[c];
        hi[i] = hi[c];
        i = c;
    }
    hk[i] = key;
    hi[i] = id;
}

static void heap_sort_out(int64_t k, float* hk, int64_t* hi) {
    /* heap_reorder: pop worst to the back until sorted best-first */
    for (int64_t n = k; n > 1; n--) {
        float key = hk[n - 1];
        int64_t id = hi[n - 1];
        hk[n - 1] = hk[0];
        hi[n - 1] = hi[0];
        heap_replace_top(n - 1, hk, hi, key, id);
    }
}

void flatref_heap_init(int64_t nq, int64_t k, float* K, int64_t* I) {
    for (int64_t i = 0; i < nq * k; i++) {
        K[i] = FLT_MAX;
        I[i] = -1;
    }
}

/* strict compare against the current k-th: an equal key does not enter */
static inline void heap_offer(int64_t k, float* hk, int64_t* hi, float key, int64_t id) {
    if (key < hk[0]) heap_replace_top(k, hk, hi, key, id);
}

static void finish(int64_t nq, int64_t k, int metric, float* K, int64_t* I) {
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < nq; q++) {
        float* hk = K + q * k;
        int64_t* hi = I + q * k;
        heap_sort_out(k, hk, hi);
        /* empty slots (id -1, key FLT_MAX) sorted last already */
        if (metric == METRIC_IP)
            for (int64_t j = 0; j < k; j++) hk[j] = (hi[j] < 0) ? -FLT_MAX : -hk[j];
    }
}

int flatref_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* nq < 20 path.  D/I are [nq, k]. */
int flatref_knn_seq(const float* xq, int64_t nq, const float* xb, int64_t n, int d, int64_t k,
                    int metric, float* D, int64_t* I) {
    if (k <= 0 || d <= 0) return 1;
    flatref_heap_init(nq, k, D, I);
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t q = 0; q < nq; q++) {
        const float* x = xq + q * (int64_t)d;
        float* hk = D + q * k;
        int64_t* hi = I + q * k;
        for (int64_t j = 0; j < n; j++) {
            const float* y = xb + j * (int64_t)d;
            float key = (metric == METRIC_L2) ? l2sqr(x, y, d) : -inner(x, y, d);
            heap_offer(k, hk, hi, key, j);
        }
    }
    finish(nq, k, metric, D, I);
    return 0;
}

/* BLAS-path handler: fold one sgemm block ip[nq_b, nb] (row stride ld) of inner
 * products for database rows j0..j0+nb into the running heaps K/I [nq_b, k].
 * L2: key = ||q||^2 + ||x||^2 - 2<q,x>, negatives clamped to 0 (qn[nq_b] and
 * bn[nb] are the squared norms of this block's queries / rows); IP: key = -ip. */
int flatref_block_update(const float* ip, int64_t nq_b, int64_t nb, int64_t ld, int64_t j0,
                         int64_t k, int metric, const float* qn, const float* bn, float* K,
                         int64_t* I) {
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < nq_b; q++) {
        const float* row = ip + q * ld;
        float* hk = K + q * k;
        int64_t* hi = I + q * k;
        if (metric == METRIC_L2) {
            const float qq = qn[q];
            for (int64_t j = 0; j < nb; j++) {
                float dis = qq + bn[j] - 2.f * row[j];
                if (dis < 0.f) dis = 0.f;
                heap_offer(k, hk, hi, dis, j0 + j);
            }
        } else {
            for (int64_t j = 0; j < nb; j++) heap_offer(k, hk, hi, -row[j], j0 + j);
        }
    }
    return 0;
}

/* sort heaps best-first and convert keys back to the metric's sign */
int flatref_heap_finish(int64_t nq, int64_t k, int metric, float* K, int64_t* I) {
    finish(nq, k, metric, K, I);
    return 0;
}

/* Pooling(mean) + Normalize; tok [B,T,H] float32, mask [B,T] int64 -> out [B,H] */
int flatref_pool_normalize(const float* tok, const int64_t* mask, int64_t B, int T, int H,
                           float* out) {
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < B; b++) {
        float* o = out + b * (int64_t)H;
        memset(o, 0, sizeof(float) * (size_t)H);
        float cnt = 0.f;
        for (int t = 0; t < T; t++) {
            float m = (float)mask[b * (int64_t)T + t];
            if (m == 0.f) continue;
            const float* r = tok + (b * (int64_t)T + t) * (int64_t)H;
            cnt += m;
#pragma omp simd
            for (int h = 0; h < H; h++) o[h] += r[h] * m;
        }
        float c = cnt < 1e-9f ? 1e-9f : cnt;
        float ss = 0.f;
        for (int h = 0; h < H; h++) {
            o[h] /= c;
            ss += o[h] * o[h];
        }
        float nr = sqrtf(ss);
        if (nr < 1e-12f) nr = 1e-12f;
        for (int h = 0; h < H; h++) o[h] /= nr;
    }
    return 0;
}

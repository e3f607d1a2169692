/*
 * bfk.h -- C ABI of libbfk.so: exact brute-force flat kNN (+ pool/normalise) on B200.
 *
 * This is the drop-in boundary for the semantic-indexer retrieval hot path of the
 * DocQA-MS reference.  The reference reaches this path through the SWIG surface of
 * the third-party faiss-cpu wheel; each entry point below names the reference call
 * site (file:line under /root/reference) and the faiss symbol it stands in for.
 * INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; bfk_last_error()
 *     then returns a thread-local message (the Python layer raises RuntimeError,
 *     mirroring faiss's FaissException -> RuntimeError).
 *   - no CPU fallback exists: without a CUDA device every compute entry point fails.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - *_is_device = 0: pointers are host memory, the call copies H2D/D2H on `stream`
 *     and returns after the results are in the host buffers.
 *     *_is_device = 1: pointers are device memory on the index's GPU; the call enqueues
 *     its work on `stream`.  add / pooling / encoder calls return without synchronising.
 *     A search through the tensor path waits for `stream` once, to read how many answers
 *     K5 could not certify (and re-runs those on the GPU) -- unless BFK_ALGO_FLAG_DEFER
 *     is set, which postpones that to bfk_search_finish; the fp32 scan never waits.
 *   - the index owns all corpus storage; inputs are borrowed for the call; outputs
 *     are caller-allocated.
 *   - every entry point that reads or changes an index takes the index's mutex for the length of the
 *     call: a search from one host thread and an add / reserve / remove_ids / reset from another are
 *     serialised, never interleaved (faiss.IndexFlat leaves that to the caller).
 *   - streams: add / add_pooled / encode_add only ENQUEUE the append on `stream`.  The library records
 *     an event behind every append and makes whatever reads the rows later on a different stream --
 *     a search, reconstruct / write_index / serialize (which copy on the legacy default stream), a
 *     regrowth copy -- wait for it first, so rows appended on a non-blocking stream are never read
 *     before they are written.  Host-pointer calls return only after their own copies have completed.
 */
#ifndef BFK_H
#define BFK_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BFK_METRIC_INNER_PRODUCT 0 /* faiss.METRIC_INNER_PRODUCT */
#define BFK_METRIC_L2 1            /* faiss.METRIC_L2 (squared distances) */

/* corpus storage: what the index keeps in HBM */
#define BFK_STORE_F32 0  /* fp32 master rows (+ bf16 scan copy + fp32 norms)           */
#define BFK_STORE_BF16 1 /* rows rounded to bf16 at add (BASELINE configs 3 and 5)     */

/* element types of encoder outputs / masks handed to the pooling entry points */
#define BFK_DT_F32 0
#define BFK_DT_F16 1
#define BFK_DT_BF16 2
#define BFK_DT_I64 3
#define BFK_DT_I32 4
#define BFK_DT_U8 5

/* search algorithm selector for bfk_search_ex */
#define BFK_ALGO_AUTO 0
#define BFK_ALGO_EXACT_SIMT 1  /* fp32 direct-distance scan (faiss's nq<20 arithmetic)  */
#define BFK_ALGO_TENSOR 2      /* bf16 tcgen05 scan + exact fp32 rerank + certificate   */

/* OR-ed into `algo`, device pointers only: enqueue the search and return without checking which
 * answers were certified exact; (D, I) are final only after bfk_search_finish.  Lets a caller queue
 * more GPU work (the row-sharded index queues its all-gather and merge) behind the scan instead of
 * idling the GPU while the host waits. */
#define BFK_ALGO_FLAG_DEFER 0x100

#define BFK_MAX_K 2048 /* faiss-cpu has no cap; faiss-gpu's historic cap is 2048 */
#define BFK_MAX_D 2048

typedef struct bfk_index bfk_index;

typedef struct bfk_search_stats {
    int32_t algo_used;        /* BFK_ALGO_* actually run for the bulk of the queries   */
    int32_t n_launches;       /* kernels launched by this call                         */
    int64_t n_fallback;       /* queries re-run exactly after a failed certificate     */
    int64_t n_candidates;     /* queries that needed the second tensor pass (tensor path) */
    float scan_ms;            /* device time of the dominant scan kernel (if timed)    */
    float reserved;
} bfk_search_stats;

const char* bfk_last_error(void);
int bfk_version(void);
/* number of CUDA devices visible; 0 (and no error) on a machine without a GPU */
int bfk_device_count(void);

/* faiss.IndexFlatL2(d) / IndexFlatIP(d) -- semantic-indexer/indexer.py:39,104 */
int bfk_create(int d, int metric, int device, int storage, bfk_index** out);
int bfk_free(bfk_index* idx);

/* index.d / index.metric_type / index.ntotal -- indexer.py:109, llm-qa/main.py:59 */
int bfk_dim(const bfk_index* idx);
int bfk_metric(const bfk_index* idx);
int bfk_storage(const bfk_index* idx);
int bfk_device(const bfk_index* idx);
int64_t bfk_ntotal(const bfk_index* idx);

/* pre-size the corpus storage (no faiss counterpart): maps device memory for n rows at once.  Growth itself
 * copies nothing -- the storage is a reserved address range with physical memory mapped behind it as rows arrive */
int bfk_reserve(bfk_index* idx, int64_t n);

/* index.add(float32[n, d]) -- indexer.py:41.  Copies; ids are sequential. */
int bfk_add(bfk_index* idx, const float* x, int64_t n, int x_is_device, void* stream);

/* SentenceTransformer.encode's Pooling(mean)+Normalize tail fused with index.add
 * -- indexer.py:37 + :41.  tok [B,T,H] (F32/F16/BF16), mask [B,T] (I64/I32/U8),
 * both device pointers; appends B unit-norm rows. */
int bfk_add_pooled(bfk_index* idx, const void* tok, int tok_dtype, const void* mask,
                   int mask_dtype, int64_t B, int T, int H, void* stream);

/* the same pooling without an index: out_f32 [B,H] -- indexer.py:37, llm-qa/main.py:25 */
int bfk_pool_normalize(const void* tok, int tok_dtype, const void* mask, int mask_dtype,
                       int64_t B, int T, int H, float* out_f32, int io_is_device, int device,
                       void* stream);

/* index.search(float32[nq, d], k) -> D float32[nq,k], I int64[nq,k]
 * -- llm-qa/main.py:101,117 (through langchain_community FAISS).
 * L2: squared distances ascending; IP: inner products descending; missing
 * results I = -1, D = +/-FLT_MAX. */
int bfk_search(bfk_index* idx, const float* q, int64_t nq, int64_t k, float* D, int64_t* I,
               int io_is_device, void* stream);
int bfk_search_ex(bfk_index* idx, const float* q, int64_t nq, int64_t k, float* D, int64_t* I,
                  int io_is_device, void* stream, int algo, int64_t id_offset,
                  bfk_search_stats* stats);

/* index.search(x, k, params=faiss.SearchParameters(sel=IDSelector...)): the same search restricted to the
 * rows whose bit is set in `bitmap` (faiss IDSelectorBitmap layout: bit i & 7 of byte i >> 3 <=> row i;
 * at least ceil(ntotal / 8) bytes; host or device memory).  Not used by the reference today; it is what the
 * patient-snippets endpoint synthese-comparative/core/retrieval_client.py:58-91 expects from the indexer
 * needs (rows of one patient / date window).  Fewer than k admitted rows -> the usual -1 padding.
 * BFK_ALGO_FLAG_DEFER is ignored. */
int bfk_search_sel(bfk_index* idx, const float* q, int64_t nq, int64_t k, float* D, int64_t* I,
                   int io_is_device, void* stream, int algo, int64_t id_offset, const uint8_t* bitmap,
                   int bitmap_is_device, bfk_search_stats* stats);

/* index.range_search(x, radius) -> (lims, D, I) -- LangChain-side extra, exact fp32
 * arithmetic: squared L2 < radius, or inner product > radius.  lims is HOST int64[nq + 1]; the (D, I) pairs
 * stay on the device behind *result until copied out (query i owns lims[i]..lims[i+1], unordered inside). */
int bfk_range_search(bfk_index* idx, const float* q, int64_t nq, float radius, int q_is_device, void* stream,
                     int64_t* lims, void** result);
int64_t bfk_range_result_size(void* result);
int bfk_range_result_copy(void* result, float* D, int64_t* I, int out_is_device);
int bfk_range_result_free(void* result);

/* settles a search started with BFK_ALGO_FLAG_DEFER: waits for it, re-runs the queries whose answer
 * was not certified (on the GPU) and reports how many result rows it rewrote; a no-op otherwise */
int bfk_search_finish(bfk_index* idx, void* stream, int64_t* n_rewritten);
/* device address of the int holding how many answers of the pending deferred search are not yet
 * certified (NULL when nothing is pending); consumed by bfk_xchg_merge */
const void* bfk_search_pending_nflag_ptr(bfk_index* idx);

/* Measurement aid (no reference counterpart; bench.py's roofline): between begin and end every tensor-path
 * pass of this index stamps the device clock (globaltimer) from inside its kernels -- when its query
 * preparation starts, when the scan's first CTA starts streaming, when the scan's last CTA leaves, when the
 * last re-rank CTA leaves -- so nothing is recorded on the stream between the kernels of a search and the
 * searches run exactly as they do unprofiled.  end waits for the device and returns, summed over the passes
 * (up to 65536 per begin; later ones are not stamped): scan kernel time; whole-pass time (preparation to
 * re-rank); device idle time between one pass's end and the next one's start (the host round trip); the
 * number of passes.  Any output pointer may be NULL. */
int bfk_profile_begin(bfk_index* idx);
int bfk_profile_end(bfk_index* idx, double* scan_ms_sum, double* pass_ms_sum, double* idle_ms_sum, int64_t* n_scans);
/* the raw stamps of the passes since the last begin, in order: out[4 * i + {0, 1, 2, 3}] = scan start, scan end,
 * preparation start, re-rank end of pass i (nanoseconds of the device clock); returns how many passes were
 * copied (at most max_passes), -1 on error */
int64_t bfk_profile_stamps(bfk_index* idx, unsigned long long* out, int64_t max_passes);

/* index.reset() / reconstruct / reconstruct_n / remove_ids -- LangChain side paths
 * on the object built at llm-qa/main.py:53-58 */
int bfk_reset(bfk_index* idx);
int bfk_reconstruct(const bfk_index* idx, int64_t i0, int64_t n, float* out, int out_is_device);
int bfk_remove_ids(bfk_index* idx, const int64_t* ids, int64_t n, int64_t* nremoved);

/* faiss.write_index / read_index -- indexer.py:27,99; llm-qa/main.py:35.
 * Byte-identical to faiss's IxF2 / IxFI files. */
int bfk_write(const bfk_index* idx, const char* path);
int bfk_read(const char* path, int device, int storage, bfk_index** out);
/* save_state() without the full rewrite -- indexer.py:26-30 calls faiss.write_index on the whole index after
 * every message (indexer.py:125).  `path` already holds the first rows_on_disk rows of this index, as a file
 * written by bfk_write / bfk_write_append / faiss; the rows after them are appended (only that tail leaves the
 * GPU) and the header is then rewritten in one 45-byte write, so the file is byte-identical to what bfk_write
 * would produce and a concurrent reader sees the old index or the new one.  Fails -- write the whole file
 * instead -- when the header does not match (other d / metric / row count, rows removed since). */
int bfk_write_append(bfk_index* idx, const char* path, int64_t rows_on_disk);
/* faiss.serialize_index / deserialize_index */
int64_t bfk_serialized_size(const bfk_index* idx);
int bfk_serialize(const bfk_index* idx, uint8_t* buf, int64_t buf_len);
int bfk_deserialize(const uint8_t* buf, int64_t len, int device, int storage, bfk_index** out);

/* faiss.normalize_L2(x): in place, host float32[n, d] (zero rows untouched) */
int bfk_normalize_l2(float* x, int64_t n, int d, int x_is_device, int device, void* stream);

/* Row-sharded search, merge step (BASELINE config 5; no reference counterpart, the
 * reference is single-process): Dparts/Iparts [nparts, nq, k] are the per-shard
 * sorted results after the all-gather; writes the global top-k.  Device pointers. */
int bfk_merge_topk(const float* Dparts, const int64_t* Iparts, int nparts, int64_t nq, int64_t k,
                   int metric, float* D, int64_t* I, int device, void* stream);
/* same, for parts that are not densely packed: part p starts at Dparts + p*d_part_stride and
 * Iparts + p*i_part_stride (strides in elements) -- the all-gather buffer can be merged in place */
int bfk_merge_topk_strided(const float* Dparts, int64_t d_part_stride, const int64_t* Iparts,
                           int64_t i_part_stride, int nparts, int64_t nq, int64_t k, int metric, float* D,
                           int64_t* I, int device, void* stream);

/* Row-sharded search, latency path (BASELINE config 5; no reference counterpart): exchange of the
 * per-GPU lists over NVLink peer memory fused with the merge, instead of a host-launched collective.
 * One process per GPU: create -> all-gather the 64-byte IPC handles by any means -> connect; then per
 * search bfk_xchg_merge(local lists) on every rank.  After synchronising `stream`,
 * bfk_xchg_last_max_nflag() > 0 (the same value on every rank) means some rank still has to settle
 * uncertified answers (bfk_search_finish) and ALL ranks must call bfk_xchg_merge again. */
int bfk_xchg_create(int device, int world, int rank, int64_t max_nq_times_k, void** out, uint8_t* ipc_handle_out);
int bfk_xchg_connect(void* handle, const uint8_t* all_handles);
int64_t bfk_xchg_capacity(void* handle);
int bfk_xchg_merge(void* handle, const float* D_local, const int64_t* I_local, int64_t nq, int64_t k, int metric,
                   const int* nflag_dev, float* D_out, int64_t* I_out, void* stream);
int64_t bfk_xchg_last_max_nflag(void* handle); /* -2: a peer's lists did not arrive within the time limit
                                                 * (60 s, BFK_XCHG_TIMEOUT_S): recreate the exchange on all ranks */
/* block_rows > 0: the ids handed to bfk_xchg_merge are LOCAL rows of a block-cyclic placement (global row g
 * lives on rank (g / block_rows) % world) and are made global while they are pushed; 0: ids are sent as they are */
int bfk_xchg_set_id_map(void* handle, int64_t block_rows);
int bfk_xchg_free(void* handle);

/* The sentence encoder feeding the pooling kernel and the index without leaving the GPU:
 * a BERT encoder of all-MiniLM-L6-v2's shape -- SentenceTransformer('all-MiniLM-L6-v2').encode at
 * semantic-indexer/indexer.py:21,37 and HuggingFaceEmbeddings at llm-qa/main.py:25, minus tokenisation.
 * This build supports hidden = 384 with 32-wide heads, intermediate % 384 == 0, T % 64 == 0, T <= 256.
 * bfk_enc_set_tensor takes host fp32 tensors under their transformers BertModel.state_dict() names
 * ("embeddings.word_embeddings.weight", "encoder.layer.0.attention.self.query.weight", ...).
 * ids int64 [B, T] and mask [B, T] (BFK_DT_I64 / I32 / U8) are DEVICE pointers; work is enqueued on stream. */
int bfk_enc_create(int device, int vocab, int hidden, int layers, int heads, int inter, int max_pos, float eps,
                   void** out);
int bfk_enc_free(void* enc);
int bfk_enc_set_tensor(void* enc, const char* name, const float* data, int64_t nelem);
/* last_hidden_state: bf16 [B, T, hidden] on the device */
int bfk_enc_forward(void* enc, const int64_t* ids, const void* mask, int mask_dtype, int64_t B, int T,
                    void* out_tok_bf16, void* stream);
/* encode(): forward + mean pool + L2 normalise -> float32 [B, hidden] on the device */
int bfk_enc_encode(void* enc, const int64_t* ids, const void* mask, int mask_dtype, int64_t B, int T, float* out_emb,
                   void* stream);
/* indexer.py:37 + :41 in one call: encode the chunks and append their embeddings to the index */
int bfk_enc_encode_add(void* enc, bfk_index* idx, const int64_t* ids, const void* mask, int mask_dtype, int64_t B,
                       int T, void* stream);

/* Exact search that allocates nothing (no counterpart in the reference, whose faiss-cpu
 * index simply mallocs).  For callers that must not call cudaMalloc once they are running:
 * processes that gave all of HBM to their own pool at start-up, request handlers that cannot
 * take the driver's allocation lock, threads that only hold pre-registered buffers.  Same
 * results as bfk_search_ex(..., BFK_ALGO_EXACT_SIMT) for nq <= 8 and k <= 128; q, D, I are
 * DEVICE pointers on the index's GPU (q row stride = d); work is only enqueued on `stream`.
 * Scratch is a fixed 1.25 MiB array inside the library image (held by every GPU that loads
 * libbfk.so until exit) shared by all calls on that GPU: calls that overlap on the GPU race;
 * calls on one stream do not.  The first call on each device resolves the array's device
 * address -- make it before any stream capture.  The first call of each kernel variant -- query
 * count rounded up to 1, 2, 4 or 8, times metric, times storage type -- may also make the driver
 * load that variant's code and set its shared-memory attribute: run every variant you will
 * serve once at start-up, and from then on a call leaves cudaMemGetInfo's free bytes untouched.
 * Larger nq / k / SM counts are rejected. */
int bfk_search_noalloc(const bfk_index* idx, const float* q_dev, int64_t nq, int64_t k, float* D_dev,
                       int64_t* I_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BFK_H */

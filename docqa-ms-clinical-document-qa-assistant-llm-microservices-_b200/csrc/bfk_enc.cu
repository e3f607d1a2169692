// The sentence encoder as an object behind the C ABI: a BERT encoder of
// all-MiniLM-L6-v2's shape (6 layers, hidden 384, 12 heads, intermediate 1536, post-LayerNorm, erf GELU)
// whose output feeds K1 (pool + normalise) and the index without leaving the GPU.
//
// Stands in for SentenceTransformer('all-MiniLM-L6-v2').encode as called at semantic-indexer/indexer.py:21,37
// and HuggingFaceEmbeddings.embed_query at llm-qa/main.py:25, minus tokenisation (host side, unchanged).
// Weights arrive under their transformers BertModel state_dict names (bfk_enc_set_tensor) as fp32 and are
// kept as bf16 [out, in] matrices (Q, K, V packed into one [1152, 384]) with fp32 biases / LayerNorm parameters.
//
// Per layer: E2 GEMM (QKV, bias) -> E3 attention -> E2 GEMM (+ residual + LayerNorm) -> E2 GEMM (GELU)
//            -> E2 GEMM (+ residual + LayerNorm); E1 embeds; K1 pools.  Chunks are processed in micro-batches
// so the activations (7.7 KB per token) stay bounded.
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "bfk_common.cuh"

namespace bfk {

enum { EPI_BIAS = 0, EPI_GELU = 1, EPI_RES_LN = 2 };
int launch_enc_gemm(int epi, const __nv_bfloat16* A, const __nv_bfloat16* W, int M, int N, int K, const float* bias,
                    const __nv_bfloat16* residual, const float* gamma, const float* beta, float eps,
                    __nv_bfloat16* out, int ldo, int num_sms, cudaStream_t st);
int launch_enc_attention(const __nv_bfloat16* qkv, const float* mask_bias, __nv_bfloat16* ctx, int64_t B, int T, int H,
                         int heads, cudaStream_t st);
int launch_enc_embed(const int64_t* ids, const void* mask, int mask_dtype, const float* word, const float* pos,
                     const float* type0, const float* gamma, const float* beta, float eps, int64_t M, int T, int H,
                     int vocab, __nv_bfloat16* x, float* mask_bias, cudaStream_t st);

struct EncLayer {
    __nv_bfloat16 *wqkv = nullptr, *wo = nullptr, *w1 = nullptr, *w2 = nullptr;
    float *bqkv = nullptr, *bo = nullptr, *b1 = nullptr, *b2 = nullptr;
    float *ln1g = nullptr, *ln1b = nullptr, *ln2g = nullptr, *ln2b = nullptr;
};

struct Encoder {
    int device = 0, num_sms = 148;
    int vocab = 0, H = 0, L = 0, heads = 0, inter = 0, max_pos = 0;
    float eps = 1e-12f;
    float *word = nullptr, *pos = nullptr, *type0 = nullptr, *emb_g = nullptr, *emb_b = nullptr;
    std::vector<EncLayer> layers;
    // activations for up to cap_rows tokens
    int64_t cap_rows = 0;
    __nv_bfloat16 *x = nullptr, *x1 = nullptr, *qkv = nullptr, *ctx = nullptr, *hid = nullptr;
    float* mask_bias = nullptr;
    std::vector<void*> owned;
    std::mutex mu;
};

static uint16_t f32_to_bf16_rn(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

template <typename T>
static int dev_alloc(Encoder* e, T** out, size_t n) {
    void* p = nullptr;
    BFK_CUDA(cudaMalloc(&p, sizeof(T) * std::max<size_t>(n, 1)));
    e->owned.push_back(p);
    *out = (T*)p;
    return 0;
}

static int upload_f32(float* dst, const float* src, size_t n) {
    BFK_CUDA(cudaMemcpy(dst, src, n * 4, cudaMemcpyHostToDevice));
    return 0;
}
static int upload_bf16(__nv_bfloat16* dst, const float* src, size_t n) {
    std::vector<uint16_t> tmp(n);
    for (size_t i = 0; i < n; i++) tmp[i] = f32_to_bf16_rn(src[i]);
    BFK_CUDA(cudaMemcpy(dst, tmp.data(), n * 2, cudaMemcpyHostToDevice));
    return 0;
}

static int ensure_rows(Encoder* e, int64_t rows) {
    if (rows <= e->cap_rows) return 0;
    BFK_CUDA(cudaDeviceSynchronize());
    void* olds[] = {e->x, e->x1, e->qkv, e->ctx, e->hid, e->mask_bias};
    for (void* p : olds)
        if (p) cudaFree(p);
    e->x = e->x1 = e->qkv = e->ctx = e->hid = nullptr;
    e->mask_bias = nullptr;
    e->cap_rows = 0;
    const size_t r = (size_t)rows;
    BFK_CUDA(cudaMalloc((void**)&e->x, r * e->H * 2));
    BFK_CUDA(cudaMalloc((void**)&e->x1, r * e->H * 2));
    BFK_CUDA(cudaMalloc((void**)&e->qkv, r * 3 * e->H * 2));
    BFK_CUDA(cudaMalloc((void**)&e->ctx, r * e->H * 2));
    BFK_CUDA(cudaMalloc((void**)&e->hid, r * e->inter * 2));
    BFK_CUDA(cudaMalloc((void**)&e->mask_bias, r * 4));
    e->cap_rows = rows;
    return 0;
}

// token states of `Bm` chunks -> e->x ([Bm*T, H] bf16)
static int forward_rows(Encoder* e, const int64_t* ids, const void* mask, int mask_dtype, int64_t Bm, int T, cudaStream_t st) {
    const int64_t M64 = Bm * T;
    if (M64 > 0x7fffffffll) return fail("encoder: micro-batch too large");
    const int M = (int)M64, H = e->H;
    BFK_TRY(launch_enc_embed(ids, mask, mask_dtype, e->word, e->pos, e->type0, e->emb_g, e->emb_b, e->eps, M, T, H,
                             e->vocab, e->x, e->mask_bias, st));
    for (const EncLayer& l : e->layers) {
        BFK_TRY(launch_enc_gemm(EPI_BIAS, e->x, l.wqkv, M, 3 * H, H, l.bqkv, nullptr, nullptr, nullptr, 0.f, e->qkv, 3 * H,
                                e->num_sms, st));
        BFK_TRY(launch_enc_attention(e->qkv, e->mask_bias, e->ctx, Bm, T, H, e->heads, st));
        BFK_TRY(launch_enc_gemm(EPI_RES_LN, e->ctx, l.wo, M, H, H, l.bo, e->x, l.ln1g, l.ln1b, e->eps, e->x1, H, e->num_sms, st));
        BFK_TRY(launch_enc_gemm(EPI_GELU, e->x1, l.w1, M, e->inter, H, l.b1, nullptr, nullptr, nullptr, 0.f, e->hid, e->inter,
                                e->num_sms, st));
        BFK_TRY(launch_enc_gemm(EPI_RES_LN, e->hid, l.w2, M, H, e->inter, l.b2, e->x1, l.ln2g, l.ln2b, e->eps, e->x, H,
                                e->num_sms, st));
    }
    return 0;
}

static int64_t micro_batch(const Encoder* e, int T) {
    (void)e;
    // tokens per pass (7.7 KB of activations each); BFK_ENC_MB_TOKENS overrides the default for experiments
    static const int64_t tokens = [] {
        const char* s = getenv("BFK_ENC_MB_TOKENS");
        const long long v = s ? atoll(s) : 0;
        return (int64_t)(v >= 256 ? v : (1 << 17));
    }();
    return std::max<int64_t>(1, tokens / T);
}

}  // namespace bfk

using namespace bfk;

extern "C" {

int bfk_enc_create(int device, int vocab, int hidden, int layers, int heads, int inter, int max_pos, float eps,
                   void** out) {
    if (!out) return fail("enc_create: null output");
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return fail("no CUDA device available: libbfk has no CPU fallback");
    }
    if (device < 0 || device >= ndev) return fail("enc_create: bad device");
    if (hidden != 384 || heads * 32 != hidden) return fail("enc_create: this build supports hidden = 384 with 32-wide heads");
    if (inter <= 0 || inter % 384 != 0 || inter % 64 != 0) return fail("enc_create: intermediate size must be a multiple of 384");
    if (vocab <= 0 || layers <= 0 || max_pos <= 0) return fail("enc_create: bad shape");
    DeviceGuard g(device);
    Encoder* e = new Encoder();
    e->device = device;
    cudaDeviceGetAttribute(&e->num_sms, cudaDevAttrMultiProcessorCount, device);
    e->vocab = vocab;
    e->H = hidden;
    e->L = layers;
    e->heads = heads;
    e->inter = inter;
    e->max_pos = max_pos;
    e->eps = eps;
    e->layers.resize((size_t)layers);
    const size_t H = (size_t)hidden, I = (size_t)inter;
    int rc = 0;
    rc |= dev_alloc(e, &e->word, (size_t)vocab * H) | dev_alloc(e, &e->pos, (size_t)max_pos * H) | dev_alloc(e, &e->type0, H) |
          dev_alloc(e, &e->emb_g, H) | dev_alloc(e, &e->emb_b, H);
    for (EncLayer& l : e->layers) {
        rc |= dev_alloc(e, &l.wqkv, 3 * H * H) | dev_alloc(e, &l.wo, H * H) | dev_alloc(e, &l.w1, I * H) | dev_alloc(e, &l.w2, H * I);
        rc |= dev_alloc(e, &l.bqkv, 3 * H) | dev_alloc(e, &l.bo, H) | dev_alloc(e, &l.b1, I) | dev_alloc(e, &l.b2, H);
        rc |= dev_alloc(e, &l.ln1g, H) | dev_alloc(e, &l.ln1b, H) | dev_alloc(e, &l.ln2g, H) | dev_alloc(e, &l.ln2b, H);
    }
    if (rc) {
        for (void* p : e->owned) cudaFree(p);
        delete e;
        return fail("enc_create: out of device memory");
    }
    *out = e;
    return 0;
}

int bfk_enc_free(void* enc) {
    if (!enc) return 0;
    Encoder* e = (Encoder*)enc;
    DeviceGuard g(e->device);
    cudaDeviceSynchronize();
    for (void* p : e->owned) cudaFree(p);
    void* act[] = {e->x, e->x1, e->qkv, e->ctx, e->hid, e->mask_bias};
    for (void* p : act)
        if (p) cudaFree(p);
    delete e;
    return 0;
}

// name = the tensor's key in transformers' BertModel.state_dict(); data = host fp32, C-contiguous
int bfk_enc_set_tensor(void* enc, const char* name, const float* data, int64_t nelem) {
    if (!enc || !name || !data) return fail("enc_set_tensor: null argument");
    Encoder* e = (Encoder*)enc;
    DeviceGuard g(e->device);
    std::lock_guard<std::mutex> lk(e->mu);
    const std::string n(name);
    const size_t H = (size_t)e->H, I = (size_t)e->inter;
    auto want = [&](size_t k) -> int {
        return (size_t)nelem == k ? 0 : fail("enc_set_tensor: " + n + " has " + std::to_string(nelem) + " elements, expected " + std::to_string(k));
    };
    if (n == "embeddings.word_embeddings.weight") { BFK_TRY(want((size_t)e->vocab * H)); return upload_f32(e->word, data, (size_t)nelem); }
    if (n == "embeddings.position_embeddings.weight") { BFK_TRY(want((size_t)e->max_pos * H)); return upload_f32(e->pos, data, (size_t)nelem); }
    if (n == "embeddings.token_type_embeddings.weight") {
        if ((size_t)nelem < H || (size_t)nelem % H) return fail("enc_set_tensor: bad token_type_embeddings size");
        return upload_f32(e->type0, data, H);  // the path only ever uses segment 0
    }
    if (n == "embeddings.LayerNorm.weight") { BFK_TRY(want(H)); return upload_f32(e->emb_g, data, H); }
    if (n == "embeddings.LayerNorm.bias") { BFK_TRY(want(H)); return upload_f32(e->emb_b, data, H); }
    const std::string pre = "encoder.layer.";
    if (n.compare(0, pre.size(), pre) == 0) {
        const size_t dot = n.find('.', pre.size());
        if (dot == std::string::npos) return fail("enc_set_tensor: unknown tensor " + n);
        const int li = atoi(n.substr(pre.size(), dot - pre.size()).c_str());
        if (li < 0 || li >= e->L) return fail("enc_set_tensor: layer index out of range in " + n);
        EncLayer& l = e->layers[(size_t)li];
        const std::string r = n.substr(dot + 1);
        if (r == "attention.self.query.weight") { BFK_TRY(want(H * H)); return upload_bf16(l.wqkv, data, H * H); }
        if (r == "attention.self.key.weight") { BFK_TRY(want(H * H)); return upload_bf16(l.wqkv + H * H, data, H * H); }
        if (r == "attention.self.value.weight") { BFK_TRY(want(H * H)); return upload_bf16(l.wqkv + 2 * H * H, data, H * H); }
        if (r == "attention.self.query.bias") { BFK_TRY(want(H)); return upload_f32(l.bqkv, data, H); }
        if (r == "attention.self.key.bias") { BFK_TRY(want(H)); return upload_f32(l.bqkv + H, data, H); }
        if (r == "attention.self.value.bias") { BFK_TRY(want(H)); return upload_f32(l.bqkv + 2 * H, data, H); }
        if (r == "attention.output.dense.weight") { BFK_TRY(want(H * H)); return upload_bf16(l.wo, data, H * H); }
        if (r == "attention.output.dense.bias") { BFK_TRY(want(H)); return upload_f32(l.bo, data, H); }
        if (r == "attention.output.LayerNorm.weight") { BFK_TRY(want(H)); return upload_f32(l.ln1g, data, H); }
        if (r == "attention.output.LayerNorm.bias") { BFK_TRY(want(H)); return upload_f32(l.ln1b, data, H); }
        if (r == "intermediate.dense.weight") { BFK_TRY(want(I * H)); return upload_bf16(l.w1, data, I * H); }
        if (r == "intermediate.dense.bias") { BFK_TRY(want(I)); return upload_f32(l.b1, data, I); }
        if (r == "output.dense.weight") { BFK_TRY(want(H * I)); return upload_bf16(l.w2, data, H * I); }
        if (r == "output.dense.bias") { BFK_TRY(want(H)); return upload_f32(l.b2, data, H); }
        if (r == "output.LayerNorm.weight") { BFK_TRY(want(H)); return upload_f32(l.ln2g, data, H); }
        if (r == "output.LayerNorm.bias") { BFK_TRY(want(H)); return upload_f32(l.ln2b, data, H); }
    }
    return fail("enc_set_tensor: unknown tensor " + n);
}

static int enc_check(Encoder* e, const void* ids, const void* mask, int mask_dtype, int64_t B, int T) {
    if (!e) return fail("null encoder");
    if (B < 0 || T <= 0) return fail("encoder: bad shape");
    if (T > e->max_pos) return fail("encoder: sequence longer than the position table");
    if (T % 64 != 0 || T > 256) return fail("encoder: pad sequences to a multiple of 64 tokens, at most 256");
    if (B > 0 && (!ids || !mask)) return fail("encoder: null input");
    if (mask_dtype != BFK_DT_I64 && mask_dtype != BFK_DT_I32 && mask_dtype != BFK_DT_U8) return fail("encoder: bad mask dtype");
    return 0;
}

static size_t mask_elem(int mask_dtype) { return mask_dtype == BFK_DT_I64 ? 8 : mask_dtype == BFK_DT_I32 ? 4 : 1; }

// token states [B, T, H] bf16 (device) for input ids int64 [B, T] + attention mask [B, T] (device)
int bfk_enc_forward(void* enc, const int64_t* ids, const void* mask, int mask_dtype, int64_t B, int T, void* out_tok_bf16,
                    void* stream) {
    Encoder* e = (Encoder*)enc;
    BFK_TRY(enc_check(e, ids, mask, mask_dtype, B, T));
    if (B == 0) return 0;
    if (!out_tok_bf16) return fail("enc_forward: null output");
    DeviceGuard g(e->device);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(e->mu);
    const int64_t mb = micro_batch(e, T);
    BFK_TRY(ensure_rows(e, std::min(B, mb) * T));
    for (int64_t b0 = 0; b0 < B; b0 += mb) {
        const int64_t Bm = std::min(mb, B - b0);
        BFK_TRY(forward_rows(e, ids + b0 * T, (const char*)mask + (size_t)b0 * T * mask_elem(mask_dtype), mask_dtype, Bm, T, st));
        BFK_CUDA(cudaMemcpyAsync((__nv_bfloat16*)out_tok_bf16 + (size_t)b0 * T * e->H, e->x, (size_t)Bm * T * e->H * 2,
                                 cudaMemcpyDeviceToDevice, st));
    }
    return 0;
}

// sentence embeddings float32 [B, H] (device): forward + K1 (mean pool over the mask, L2 normalise)
int bfk_enc_encode(void* enc, const int64_t* ids, const void* mask, int mask_dtype, int64_t B, int T, float* out_emb,
                   void* stream) {
    Encoder* e = (Encoder*)enc;
    BFK_TRY(enc_check(e, ids, mask, mask_dtype, B, T));
    if (B == 0) return 0;
    if (!out_emb) return fail("enc_encode: null output");
    DeviceGuard g(e->device);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(e->mu);
    const int64_t mb = micro_batch(e, T);
    BFK_TRY(ensure_rows(e, std::min(B, mb) * T));
    for (int64_t b0 = 0; b0 < B; b0 += mb) {
        const int64_t Bm = std::min(mb, B - b0);
        const void* mk = (const char*)mask + (size_t)b0 * T * mask_elem(mask_dtype);
        BFK_TRY(forward_rows(e, ids + b0 * T, mk, mask_dtype, Bm, T, st));
        BFK_TRY(launch_pool(e->x, BFK_DT_BF16, mk, mask_dtype, Bm, T, e->H, out_emb + (size_t)b0 * e->H, e->H, nullptr, 0,
                            nullptr, nullptr, st));
    }
    return 0;
}

// indexer.py:37 + :41 in one call: encode the chunks and append their embeddings to the index
int bfk_enc_encode_add(void* enc, bfk_index* idx, const int64_t* ids, const void* mask, int mask_dtype, int64_t B, int T,
                       void* stream) {
    Encoder* e = (Encoder*)enc;
    BFK_TRY(enc_check(e, ids, mask, mask_dtype, B, T));
    if (!idx) return fail("null index");
    if (bfk_dim(idx) != e->H) return fail("enc_encode_add: index dimension != encoder width");
    if (bfk_device(idx) != e->device) return fail("enc_encode_add: index and encoder live on different GPUs");
    if (B == 0) return 0;
    DeviceGuard g(e->device);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(e->mu);
    const int64_t mb = micro_batch(e, T);
    BFK_TRY(ensure_rows(e, std::min(B, mb) * T));
    BFK_TRY(grow_for(idx, bfk_ntotal(idx) + B, st));  // once for all micro-batches, amortised doubling
    for (int64_t b0 = 0; b0 < B; b0 += mb) {
        const int64_t Bm = std::min(mb, B - b0);
        const void* mk = (const char*)mask + (size_t)b0 * T * mask_elem(mask_dtype);
        BFK_TRY(forward_rows(e, ids + b0 * T, mk, mask_dtype, Bm, T, st));
        BFK_TRY(bfk_add_pooled(idx, e->x, BFK_DT_BF16, mk, mask_dtype, Bm, T, e->H, stream));
    }
    return 0;
}

}  // extern "C"

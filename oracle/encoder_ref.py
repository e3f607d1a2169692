"""CPU restatement of the sentence encoder (TEST INFRASTRUCTURE ONLY: tests/, smoke() and bench.py's CPU legs).

The arithmetic lives in third-party wheels the reference does not pin (sentence-transformers -> transformers
BertModel; semantic-indexer/requirements.txt:1), so this restates the published algorithm
(transformers/models/bert/modeling_bert.py: BertEmbeddings, BertSelfAttention, BertSelfOutput,
BertIntermediate, BertOutput) in numpy float32 and is pinned by tests/test_oracle_encoder.py against
transformers' own BertModel, which IS importable in this image (random-initialised weights of
all-MiniLM-L6-v2's shape; the checkpoint itself is not available offline).

    emb   = LayerNorm(word[ids] + pos[0..T) + type[0])
    layer: q,k,v = x Wq^T + bq, ...; per head softmax(q k^T / sqrt(dh) + (1 - mask) * finfo.min) v
           x = LayerNorm(x + ctx Wo^T + bo);  x = LayerNorm(x + gelu_erf(x W1^T + b1) W2^T + b2)
"""
from __future__ import annotations

import math

import numpy as np

try:
    from scipy.special import erf as _erf
except Exception:  # pragma: no cover - scipy is in the image; the fallback is just slow
    _erf = np.vectorize(math.erf, otypes=[np.float32])


def _ln(x, g, b, eps):
    mu = x.mean(-1, keepdims=True)
    var = ((x - mu) ** 2).mean(-1, keepdims=True)
    return ((x - mu) / np.sqrt(var + eps) * g + b).astype(np.float32)


def _gelu(x):
    return (0.5 * x * (1.0 + _erf(x / np.float32(math.sqrt(2.0))))).astype(np.float32)


def forward(sd: dict, input_ids: np.ndarray, attention_mask: np.ndarray, *, num_heads: int = 12,
            eps: float = 1e-12) -> np.ndarray:
    """last_hidden_state float32 [B, T, H]; `sd` maps BertModel.state_dict() names to float32 arrays."""
    W = {k: np.asarray(v, dtype=np.float32) for k, v in sd.items()}
    ids = np.asarray(input_ids)
    B, T = ids.shape
    H = W["embeddings.word_embeddings.weight"].shape[1]
    dh = H // num_heads
    x = (W["embeddings.word_embeddings.weight"][ids] + W["embeddings.position_embeddings.weight"][None, :T]
         + W["embeddings.token_type_embeddings.weight"][0][None, None])
    x = _ln(x, W["embeddings.LayerNorm.weight"], W["embeddings.LayerNorm.bias"], eps)
    bias = ((1.0 - np.asarray(attention_mask, dtype=np.float32)) * np.finfo(np.float32).min)[:, None, None, :]
    li = 0
    while f"encoder.layer.{li}.attention.self.query.weight" in W:
        p = f"encoder.layer.{li}."

        def lin(a, name):
            return (a @ W[p + name + ".weight"].T + W[p + name + ".bias"]).astype(np.float32)

        def heads(a):
            return a.reshape(B, T, num_heads, dh).transpose(0, 2, 1, 3)

        q, k, v = heads(lin(x, "attention.self.query")), heads(lin(x, "attention.self.key")), heads(lin(x, "attention.self.value"))
        s = (q @ k.transpose(0, 1, 3, 2)) / np.float32(math.sqrt(dh)) + bias
        s = s - s.max(-1, keepdims=True)
        e = np.exp(s)
        pr = (e / e.sum(-1, keepdims=True)).astype(np.float32)
        ctx = (pr @ v).transpose(0, 2, 1, 3).reshape(B, T, H)
        x = _ln(x + lin(ctx, "attention.output.dense"), W[p + "attention.output.LayerNorm.weight"],
                W[p + "attention.output.LayerNorm.bias"], eps)
        h = _gelu(lin(x, "intermediate.dense"))
        x = _ln(x + lin(h, "output.dense"), W[p + "output.LayerNorm.weight"], W[p + "output.LayerNorm.bias"], eps)
        li += 1
    return x


def encode(sd: dict, input_ids, attention_mask, **kw) -> np.ndarray:
    """SentenceTransformer.encode on tokenised input: forward + Pooling(mean) + Normalize -> float32 [B, H]."""
    from oracle import pool_ref

    tok = forward(sd, input_ids, attention_mask, **kw)
    return pool_ref.mean_pool_normalize(tok, np.asarray(attention_mask))

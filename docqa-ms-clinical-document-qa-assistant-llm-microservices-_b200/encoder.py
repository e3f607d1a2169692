"""The sentence encoder on the GPU, feeding the pooling kernel and the index directly.

Host mirror of what `SentenceTransformer('all-MiniLM-L6-v2')` is to semantic-indexer/indexer.py:21,37 and
`HuggingFaceEmbeddings` to llm-qa/main.py:25: a BERT encoder (6 layers, hidden 384, 12 heads, intermediate
1536) followed by Pooling(mean) + Normalize.  Tokenisation stays on the host and is not part of this module:
callers pass `input_ids` / `attention_mask` (the tokenizer's output), padded to a multiple of 64 tokens.

    enc = MiniLMEncoder.from_state_dict(bert_model.state_dict())      # transformers BertModel names
    tok = enc.forward(input_ids, attention_mask)                      # last_hidden_state, bf16 CUDA [B, T, 384]
    emb = enc.encode_ids(input_ids, attention_mask)                   # float32 CUDA [B, 384], unit norm
    enc.encode_add(index, input_ids, attention_mask)                  # indexer.py:37 + :41 on the device

`encode_tokens(texts)` (what `indexer_host.SemanticIndexer.add_many` looks for) is available when a tokenizer
callable is supplied.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from .pooling import _mask_dtype

MINILM_L6 = dict(vocab_size=30522, hidden_size=384, num_hidden_layers=6, num_attention_heads=12,
                 intermediate_size=1536, max_position_embeddings=512, layer_norm_eps=1e-12)


def pad_to_multiple(input_ids, attention_mask, multiple: int = 64):
    """Right-pad [B, T] id / mask tensors with zeros so that T is a multiple of `multiple` (mask 0 = ignored)."""
    import torch

    T = input_ids.shape[1]
    Tp = -(-T // multiple) * multiple
    if Tp == T:
        return input_ids, attention_mask
    pad = (0, Tp - T)
    return torch.nn.functional.pad(input_ids, pad), torch.nn.functional.pad(attention_mask, pad)


class MiniLMEncoder:
    def __init__(self, *, device: int = 0, tokenizer=None, **config):
        cfg = {**MINILM_L6, **config}
        self.config = cfg
        self.device = int(device)
        self.tokenizer = tokenizer
        self._L = _lib.load()
        h = ctypes.c_void_p()
        _lib.check(self._L.bfk_enc_create(self.device, cfg["vocab_size"], cfg["hidden_size"], cfg["num_hidden_layers"],
                                           cfg["num_attention_heads"], cfg["intermediate_size"],
                                           cfg["max_position_embeddings"], float(cfg["layer_norm_eps"]), ctypes.byref(h)))
        self._h = h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._L.bfk_enc_free(h)
            except Exception:
                pass

    @property
    def hidden_size(self) -> int:
        return self.config["hidden_size"]

    # ---- weights ----------------------------------------------------------------------------
    def load_state_dict(self, state_dict) -> list[str]:
        """Load tensors named as in transformers' BertModel.state_dict(); returns the keys that were ignored
        (pooler.*, position_ids and anything else the sentence-embedding path never touches)."""
        ignored = []
        for name, t in state_dict.items():
            key = name[5:] if name.startswith("bert.") else name
            if key.startswith("pooler.") or key.endswith("position_ids") or key.endswith("token_type_ids"):
                ignored.append(name)
                continue
            a = t.detach().to("cpu").float().numpy() if hasattr(t, "detach") else np.asarray(t, dtype=np.float32)
            a = np.ascontiguousarray(a, dtype=np.float32)
            _lib.check(self._L.bfk_enc_set_tensor(self._h, key.encode(), ctypes.c_void_p(a.ctypes.data), a.size))
        return ignored

    @classmethod
    def from_state_dict(cls, state_dict, *, device: int = 0, tokenizer=None, **config):
        enc = cls(device=device, tokenizer=tokenizer, **config)
        enc.load_state_dict(state_dict)
        return enc

    @classmethod
    def from_hf_model(cls, model, *, device: int = 0, tokenizer=None):
        c = model.config
        return cls.from_state_dict(model.state_dict(), device=device, tokenizer=tokenizer, vocab_size=c.vocab_size,
                                   hidden_size=c.hidden_size, num_hidden_layers=c.num_hidden_layers,
                                   num_attention_heads=c.num_attention_heads, intermediate_size=c.intermediate_size,
                                   max_position_embeddings=c.max_position_embeddings, layer_norm_eps=c.layer_norm_eps)

    # ---- compute ------------------------------------------------------------------------------
    def _inputs(self, input_ids, attention_mask):
        import torch

        assert input_ids.is_cuda and attention_mask.is_cuda and input_ids.device.index == self.device
        assert input_ids.shape == attention_mask.shape and input_ids.dim() == 2
        ids = input_ids.to(torch.int64).contiguous()
        mask = attention_mask.contiguous()
        st = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        return ids, mask, st

    def forward(self, input_ids, attention_mask):
        """last_hidden_state as bf16 CUDA tensor [B, T, hidden]."""
        import torch

        ids, mask, st = self._inputs(input_ids, attention_mask)
        B, T = ids.shape
        out = torch.empty((B, T, self.hidden_size), dtype=torch.bfloat16, device=ids.device)
        _lib.check(self._L.bfk_enc_forward(self._h, ctypes.c_void_p(ids.data_ptr()), ctypes.c_void_p(mask.data_ptr()),
                                            _mask_dtype(mask), B, T, ctypes.c_void_p(out.data_ptr()), st))
        return out

    def encode_ids(self, input_ids, attention_mask):
        """Sentence embeddings float32 CUDA [B, hidden]: forward + mean pool + L2 normalise."""
        import torch

        ids, mask, st = self._inputs(input_ids, attention_mask)
        B, T = ids.shape
        out = torch.empty((B, self.hidden_size), dtype=torch.float32, device=ids.device)
        _lib.check(self._L.bfk_enc_encode(self._h, ctypes.c_void_p(ids.data_ptr()), ctypes.c_void_p(mask.data_ptr()),
                                           _mask_dtype(mask), B, T, ctypes.c_void_p(out.data_ptr()), st))
        return out

    def encode_add(self, index, input_ids, attention_mask) -> None:
        """Encode the chunks and append their embeddings to `index` (an IndexFlat on the same GPU)."""
        ids, mask, st = self._inputs(input_ids, attention_mask)
        B, T = ids.shape
        _lib.check(self._L.bfk_enc_encode_add(self._h, index._h, ctypes.c_void_p(ids.data_ptr()),
                                               ctypes.c_void_p(mask.data_ptr()), _mask_dtype(mask), B, T, st))

    # ---- SentenceTransformer-shaped entry points (need a tokenizer callable) --------------------
    def _tokenize(self, texts):
        import torch

        if self.tokenizer is None:
            raise RuntimeError("MiniLMEncoder: no tokenizer was supplied (tokenisation is host-side and not part of libbfk)")
        enc = self.tokenizer(list(texts), padding=True, truncation=True, max_length=256, return_tensors="pt")
        dev = torch.device("cuda", self.device)
        return pad_to_multiple(enc["input_ids"].to(dev), enc["attention_mask"].to(dev))

    def encode_tokens(self, texts):
        """(token_embeddings [B, T, H], attention_mask [B, T]) on the device, for IndexFlat.add_pooled."""
        ids, mask = self._tokenize(texts)
        return self.forward(ids, mask), mask

    def encode(self, texts):
        """SentenceTransformer.encode(list[str]) -> float32 numpy [n, hidden] (indexer.py:37: the text as it is)."""
        single = isinstance(texts, str)
        ids, mask = self._tokenize([texts] if single else texts)
        out = self.encode_ids(ids, mask).cpu().numpy()
        return out[0] if single else out

    # ---- langchain HuggingFaceEmbeddings-shaped entry points (llm-qa/main.py:25) -------------------
    # HuggingFaceEmbeddings replaces every newline by a space before it encodes, on the document side and on the
    # query side alike; llm-qa embeds its questions through it, so a question with line breaks
    # must be embedded the same way here or it lands somewhere else in the index.
    def embed_documents(self, texts) -> list[list[float]]:
        texts = [t.replace("\n", " ") for t in texts]
        return self.encode(texts).tolist() if texts else []

    def embed_query(self, text: str) -> list[float]:
        return self.embed_documents([text])[0]

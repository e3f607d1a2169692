"""Pins the numpy restatement of the encoder (oracle/encoder_ref.py) to transformers' BertModel. CPU only."""
import numpy as np
import pytest

from oracle import encoder_ref


def hf_model(layers=2, vocab=500, seed=0):
    torch = pytest.importorskip("torch")
    transformers = pytest.importorskip("transformers")
    cfg = transformers.BertConfig(vocab_size=vocab, hidden_size=384, num_hidden_layers=layers, num_attention_heads=12,
                                  intermediate_size=1536, max_position_embeddings=512)
    torch.manual_seed(seed)
    m = transformers.BertModel(cfg, add_pooling_layer=False).eval()
    # default init (std 0.02, LayerNorm 1/0, zero biases) exercises little: perturb biases and LayerNorm too
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif "LayerNorm.weight" in name:
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            elif name.endswith("dense.weight") or "self." in name:
                p.copy_(0.05 * torch.randn(p.shape, generator=g))
    return m


def inputs(B, T, vocab, seed=3):
    rng = np.random.default_rng(seed)
    ids = rng.integers(0, vocab, size=(B, T))
    mask = np.ones((B, T), dtype=np.int64)
    for b in range(1, B):
        mask[b, rng.integers(4, T):] = 0
    return ids, mask


def test_oracle_matches_transformers_bert():
    import torch

    m = hf_model()
    sd = {k: v.numpy() for k, v in m.state_dict().items()}
    ids, mask = inputs(3, 64, 500)
    with torch.no_grad():
        want = m(input_ids=torch.from_numpy(ids), attention_mask=torch.from_numpy(mask)).last_hidden_state.numpy()
    got = encoder_ref.forward(sd, ids, mask)
    assert got.shape == want.shape == (3, 64, 384)
    np.testing.assert_allclose(got, want, atol=2e-4, rtol=1e-4)
    # sentence embedding = Pooling(mean) + Normalize of those states, padding ignored
    emb = encoder_ref.encode(sd, ids, mask)
    t = torch.from_numpy(want)
    mk = torch.from_numpy(mask).unsqueeze(-1).float()
    ref = (t * mk).sum(1) / mk.sum(1).clamp(min=1e-9)
    ref = torch.nn.functional.normalize(ref, p=2, dim=1).numpy()
    np.testing.assert_allclose(emb, ref, atol=1e-5)

"""K1 pool+normalise(+add) against the oracle. GPU only."""
import numpy as np
import pytest

from oracle import pool_ref

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def bfk():
    import docqa_b200

    assert docqa_b200.get_num_gpus() >= 1
    return docqa_b200


def make(rng, B, T, H, lengths=None):
    tok = rng.standard_normal((B, T, H)).astype(np.float32)
    if lengths is None:
        lengths = rng.integers(1, T + 1, size=B)
    mask = (np.arange(T)[None, :] < np.asarray(lengths)[:, None]).astype(np.int64)
    return tok, mask


@pytest.mark.parametrize("shape", [(1, 256, 384), (7, 256, 384), (5, 33, 100), (3, 1, 7), (2, 512, 768), (4, 16, 2048)])
def test_pool_numpy_fp32(bfk, shape):
    rng = np.random.default_rng(sum(shape))
    tok, mask = make(rng, *shape)
    out = bfk.pool_normalize(tok, mask)
    ref = pool_ref.mean_pool_normalize_f64(tok, mask)
    assert out.dtype == np.float32 and out.shape == (shape[0], shape[2])
    np.testing.assert_allclose(out, ref, rtol=1e-5, atol=2e-6)  # fp32 summation-order tolerance


def test_pool_edge_masks(bfk):
    rng = np.random.default_rng(1)
    tok, mask = make(rng, 4, 64, 384, lengths=[64, 1, 0, 17])
    mask[3, ::2] = 0  # holes, not just a prefix
    out = bfk.pool_normalize(tok, mask)
    ref = pool_ref.mean_pool_normalize_f64(tok, mask)
    np.testing.assert_allclose(out, ref, rtol=1e-5, atol=2e-6)
    assert (out[2] == 0).all()  # all-masked row -> clamps give zeros, not NaN


@pytest.mark.parametrize("tdt", ["float32", "bfloat16", "float16"])
@pytest.mark.parametrize("mdt", ["int64", "int32", "bool"])
def test_pool_torch_dtypes_and_add_pooled(bfk, tdt, mdt):
    import torch

    rng = np.random.default_rng(2)
    B, T, H = 37, 256, 384
    tok, mask = make(rng, B, T, H)
    t = torch.from_numpy(tok).cuda().to(getattr(torch, tdt))
    m = torch.from_numpy(mask).cuda().to(getattr(torch, mdt))
    ref = pool_ref.mean_pool_normalize_f64(t.float().cpu().numpy(), mask)
    out = bfk.pool_normalize(t, m)
    assert out.is_cuda and out.dtype == torch.float32
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=1e-5, atol=2e-6)
    # unaligned view -> scalar-load path, same answer
    buf = torch.empty(B * T * H + 1, dtype=t.dtype, device="cuda")
    t2 = buf[1:].view(B, T, H)
    t2.copy_(t)
    out2 = bfk.pool_normalize(t2, m)
    np.testing.assert_allclose(out2.cpu().numpy(), ref, rtol=1e-5, atol=2e-6)
    # fused append: rows land in the index exactly as pool_normalize returns them
    idx = bfk.IndexFlatL2(H)
    idx.add(out[:5])
    idx.add_pooled(t, m)
    assert idx.ntotal == 5 + B
    got = idx.reconstruct_n(5, B)
    assert (got == out.cpu().numpy()).all()
    D, I = idx.search(out[3:4], 2)
    assert sorted(I.cpu().numpy()[0].tolist()) == [3, 8] and float(D.max()) < 1e-10

"""Pooling oracle: numpy == C == float64 formulas, incl. clamps (not gpu)."""
import numpy as np

from oracle import cpu_baseline, pool_ref


def test_pool_formulas_and_clamps():
    rng = np.random.default_rng(0)
    B, T, H = 6, 17, 24
    tok = rng.standard_normal((B, T, H)).astype(np.float32)
    mask = np.ones((B, T), np.int64)
    mask[1, 5:] = 0
    mask[2, :] = 0  # all-masked row: sum_mask clamps to 1e-9, norm clamps to 1e-12 -> zeros
    mask[3, 1:] = 0
    out = pool_ref.mean_pool_normalize(tok, mask)
    ref = pool_ref.mean_pool_normalize_f64(tok, mask)
    assert out.dtype == np.float32 and out.shape == (B, H)
    np.testing.assert_allclose(out, ref, rtol=2e-6, atol=2e-7)
    assert (out[2] == 0).all()
    np.testing.assert_allclose(np.linalg.norm(out[[0, 1, 3, 4, 5]], axis=1), 1, atol=1e-6)
    # single valid token -> that token, normalised
    np.testing.assert_allclose(out[3], tok[3, 0] / np.linalg.norm(tok[3, 0]), rtol=1e-6, atol=1e-7)
    c = cpu_baseline.pool_normalize(tok, mask)
    np.testing.assert_allclose(c, ref, rtol=2e-6, atol=2e-7)


def test_pool_outputs_look_like_the_reference_index(fixture_corpus):
    # the only numeric pin the reference offers for this stage: shipped rows are unit norm
    rng = np.random.default_rng(1)
    tok = rng.standard_normal((4, 256, 384)).astype(np.float32)
    mask = (np.arange(256)[None, :] < np.array([256, 16, 100, 1])[:, None]).astype(np.int64)
    out = pool_ref.mean_pool_normalize(tok, mask)
    a = np.linalg.norm(out.astype(np.float64), axis=1)
    b = np.linalg.norm(fixture_corpus.astype(np.float64), axis=1)
    assert np.abs(a - 1).max() < 3e-7 and np.abs(b - 1).max() < 3e-7

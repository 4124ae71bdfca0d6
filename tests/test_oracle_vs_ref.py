"""Pins the C restatement (oracle/lattice_oracle.c) against the reference itself
(oracle/_ref, the reference header compiled in place).  CPU only."""
import numpy as np
import pytest

import importlib
synthetic = importlib.import_module("depth_estimation_b200.synthetic")

CASES = [  # n, d, C, feature scale
    (2000, 5, 3, 10.0), (5000, 3, 6, 5.0), (3000, 2, 4, 20.0), (1500, 10, 2, 2.0),
    (4000, 5, 7, 100.0), (100, 1, 2, 3.0), (80, 3, 2, 1.0), (1, 5, 1, 1.0),
]


def _bits(a):
    return a.view(np.uint32) if a.dtype == np.float32 else a


def _rand(n, d, C, scale, seed=0):
    rng = np.random.default_rng(seed)
    ref = (rng.random((n, d), dtype=np.float32) * scale - scale / 3).astype(np.float32)
    src = rng.standard_normal((n, C)).astype(np.float32)
    return src, ref


@pytest.mark.parametrize("n,d,C,scale", CASES)
def test_faithful_oracle_is_bit_exact(oracle_faithful, reference, n, d, C, scale):
    src, ref = _rand(n, d, C, scale)
    so, sr = oracle_faithful.state(ref), reference.state(ref)
    assert so.M == sr.M
    for k in ("greedy", "rank", "bary", "vid", "keys", "nbr"):
        a, b = getattr(so, k), getattr(sr, k)
        assert a.shape == b.shape, k
        assert np.array_equal(_bits(a), _bits(b)), k
    assert np.array_equal(_bits(oracle_faithful.filter(src, ref)), _bits(reference.filter(src, ref)))


@pytest.mark.parametrize("n,d,scale", [(60000, 5, 6.0), (120000, 5, 12.0)])
def test_faithful_oracle_reproduces_grow_quirk(oracle_faithful, oracle_clean, reference, n, d, scale):
    """Past 16383 vertices the reference's table doubles and its stale-hash lookup
    (permutohedral.h:102-103,62) can add orphan duplicates; the faithful oracle
    must follow it exactly, the clean oracle must equal its distinct-key set."""
    src, ref = _rand(n, d, 2, scale, seed=3)
    so, sr, sc = oracle_faithful.state(ref), reference.state(ref), oracle_clean.state(ref)
    assert so.M == sr.M and np.array_equal(so.vid, sr.vid) and np.array_equal(so.keys, sr.keys)
    assert np.array_equal(so.nbr, sr.nbr)
    assert np.array_equal(_bits(oracle_faithful.filter(src, ref)), _bits(reference.filter(src, ref)))
    uniq = np.unique(sr.keys, axis=0)
    assert sc.M == len(uniq)
    assert 0 <= sr.M - sc.M <= 8  # at most one orphan per doubling
    assert np.array_equal(np.unique(sc.keys, axis=0), uniq)
    assert np.array_equal(so.greedy, sc.greedy) and np.array_equal(so.rank, sc.rank)


def test_clean_equals_faithful_without_growth(oracle_faithful, oracle_clean):
    img = synthetic.image(96, 128, seed=1)
    ref = synthetic.ijrgb_features(img)
    rng = np.random.default_rng(0)
    src = rng.random((ref.shape[0], 5)).astype(np.float32)
    a, b = oracle_clean.build(ref), oracle_faithful.build(ref)
    assert a.ngrow == 0 and a.M == b.M
    assert np.array_equal(_bits(a.apply(src)), _bits(b.apply(src)))


def test_scale_factor_quirk_and_constants(oracle):
    # permutohedral.h:356,370: fp32 expression, not the exact real value
    sf = oracle.scale_factors(5)
    want = np.array([(1.0 / np.sqrt(np.float32((i + 1) * (i + 2)), dtype=np.float32))
                     for i in range(5)], np.float32)
    want = (want * np.float32(6 * np.sqrt(np.float32(2.0 / 3), dtype=np.float32))).astype(np.float32)
    assert np.allclose(sf, want, rtol=1e-6)


def test_filter_of_ones_is_positive_degree(oracle):
    # RbfLaplacian's D = filter(1) (gaussian_matrix.py:314) must be > 0
    img = synthetic.image(48, 64, seed=2)
    ref = synthetic.ijrgb_features(img)
    D = oracle.filter(np.ones((ref.shape[0], 1), np.float32), ref)
    assert (D > 0).all()


def test_filter_is_linear_and_nearly_symmetric(oracle):
    rng = np.random.default_rng(5)
    ref = (rng.random((600, 3), dtype=np.float32) * 4).astype(np.float32)
    h = oracle.build(ref)
    a = rng.standard_normal((600, 2)).astype(np.float32)
    b = rng.standard_normal((600, 2)).astype(np.float32)
    assert np.allclose(h.apply(a + 2 * b), h.apply(a) + 2 * h.apply(b), atol=2e-4)
    # <a, K b> ~ <K a, b>: blur passes commute only approximately
    lhs, rhs = (a * h.apply(b)).sum(), (h.apply(a) * b).sum()
    assert abs(lhs - rhs) <= 0.3 * max(abs(lhs), abs(rhs), 1.0)

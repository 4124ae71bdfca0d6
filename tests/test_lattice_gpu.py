"""GPU parity: CUDA lattice (through the C ABI) vs the CPU oracle on the same inputs.

Integer artefacts (greedy, rank, keys, vertex count, vertex ids, neighbour
indices) must be bit-exact; barycentric weights bit-exact (same fp32 op order);
filtered values within 1e-4 of the output scale (atomic-order nondeterminism).
"""
import importlib

import numpy as np
import pytest
import torch

synthetic = importlib.import_module("depth_estimation_b200.synthetic")

pytestmark = pytest.mark.gpu

TOL = 1e-4  # max |gpu - oracle| / max|oracle|   (north_star: max rel err 1e-4)


def _rand(n, d, C, scale, seed=0):
    rng = np.random.default_rng(seed)
    ref = (rng.random((n, d), dtype=np.float32) * scale - scale / 3).astype(np.float32)
    src = rng.standard_normal((n, C)).astype(np.float32)
    return src, ref


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _check_state(lat, st):
    ex = {k: v.cpu().numpy() for k, v in lat.export().items()}
    assert lat.num_vertices == st.M
    assert np.array_equal(ex["greedy"], st.greedy)
    assert np.array_equal(ex["rank"], st.rank)
    assert np.array_equal(ex["bary"].view(np.uint32), st.bary.view(np.uint32))
    assert np.array_equal(ex["keys"], st.keys)
    assert np.array_equal(ex["vid"], st.vid)
    assert np.array_equal(ex["nbr"], st.nbr)


def _relerr(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("n,d,C,scale", [
    (2000, 5, 3, 10.0), (5000, 3, 6, 5.0), (3000, 2, 4, 20.0), (1500, 10, 2, 2.0),
    (100, 1, 2, 3.0), (80, 3, 2, 1.0), (1, 5, 1, 1.0), (33, 5, 8, 4.0), (4097, 4, 5, 7.0),
    (2500, 16, 4, 1.0), (60000, 5, 2, 6.0),
])
def test_random_features_match_oracle(native, oracle, n, d, C, scale):
    src, ref = _rand(n, d, C, scale)
    st = oracle.state(ref)
    lat = native.PermutohedralLattice(_cuda(ref))
    _check_state(lat, st)
    want = oracle.filter(src, ref)
    got = lat.filter(_cuda(src)).cpu().numpy()
    assert _relerr(got, want) < TOL
    got_exact = lat.filter(_cuda(src), exact_slice=True).cpu().numpy()
    assert _relerr(got_exact, want) < TOL
    # flat per-entry kernels (A/B switch) agree with the tile-local ones
    got_flat = lat.filter(_cuda(src), flat=True).cpu().numpy()
    assert _relerr(got_flat, want) < TOL
    # one-shot entry point with the reference's call shape
    got2 = native.lattice.filter(_cuda(src), _cuda(ref)).cpu().numpy()
    assert _relerr(got2, want) < TOL


@pytest.mark.parametrize("h,w,s_ij,s_rgb,C", [
    (96, 128, 0.1, 0.1, 16), (288, 384, 0.1, 0.1, 16), (288, 384, 0.01, 0.125, 4),
    (120, 160, 0.02, 0.05, 64),
])
def test_image_features_match_oracle(native, oracle, h, w, s_ij, s_rgb, C):
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=4), s_ij, s_rgb)
    rng = np.random.default_rng(1)
    src = rng.random((h * w, C)).astype(np.float32)
    st = oracle.state(ref)
    lat = native.PermutohedralLattice(_cuda(ref))
    _check_state(lat, st)
    want = oracle.build(ref).apply(src)
    got = lat.filter(_cuda(src)).cpu().numpy()
    assert _relerr(got, want) < TOL
    assert _relerr(lat.filter(_cuda(src), flat=True).cpu().numpy(), want) < TOL
    assert _relerr(lat.filter(_cuda(src), exact_slice=True).cpu().numpy(), want) < TOL


def test_matches_reference_itself_when_its_table_never_grows(native, reference):
    """M < 16383: the reference's hash table never doubles, so the unmodified
    reference (oracle/_ref) is directly comparable, vertex ids included."""
    ref = synthetic.ijrgb_features(synthetic.image(288, 384, seed=0))  # Tsukuba size
    rng = np.random.default_rng(2)
    src = rng.random((ref.shape[0], 16)).astype(np.float32)
    st = reference.state(ref)
    assert st.M < 16383
    lat = native.PermutohedralLattice(_cuda(ref))
    _check_state(lat, st)
    assert _relerr(lat.filter(_cuda(src)).cpu().numpy(), reference.filter(src, ref)) < TOL


def test_host_buffers_roundtrip(native, oracle):
    """lattice.filter on CPU tensors: H2D, filter, D2H inside the call (the reference's call shape)."""
    src, ref = _rand(3000, 5, 4, 8.0, seed=7)
    out = native.lattice.filter(torch.from_numpy(src), torch.from_numpy(ref))
    assert not out.is_cuda
    assert _relerr(out.numpy(), oracle.filter(src, ref)) < TOL


def test_fused_epilogue_and_strided_channels(native, oracle):
    src, ref = _rand(4000, 5, 12, 6.0, seed=8)
    rng = np.random.default_rng(9)
    add = rng.standard_normal(src.shape).astype(np.float32)
    lat = native.PermutohedralLattice(_cuda(ref))
    want = oracle.filter(src, ref)
    s = _cuda(src)
    got = lat.filter(s, sub=s, add=_cuda(add)).cpu().numpy()
    assert _relerr(got, add + (want - src)) < TOL
    # a column block of a wider matrix (row stride > C), as the backward uses
    wide = torch.zeros((4000, 20), device="cuda")
    wide[:, 4:16] = s
    got = lat.filter(wide[:, 4:16]).cpu().numpy()
    assert _relerr(got, want) < TOL
    # odd channel count -> scalar path
    got = lat.filter(s[:, :7].contiguous()).cpu().numpy()
    assert _relerr(got, want[:, :7]) < TOL


@pytest.mark.parametrize("n,d,C,scale", [
    (3000, 5, 40, 4.0), (3000, 5, 64, 0.5), (2500, 5, 100, 12.0), (2000, 3, 48, 3.0), (2000, 2, 72, 25.0),
    (1500, 7, 36, 2.0), (900, 10, 64, 1.0), (130, 5, 256, 2.0), (5000, 5, 68, 30.0),
])
def test_wide_channels_use_the_tiled_kernels(native, oracle, n, d, C, scale):
    """C > 32 takes the tile-local splat / slice (pipelined splat for d = 5, range splat otherwise;
    slice staging 48 / 96 / 160 rows or unstaged tiles at large feature scales; partial last channel
    block), with and without the fused epilogue."""
    src, ref = _rand(n, d, C, scale, seed=n + d + C)
    lat = native.PermutohedralLattice(_cuda(ref))
    want = oracle.filter(src, ref)
    s = _cuda(src)
    assert _relerr(lat.filter(s).cpu().numpy(), want) < TOL
    assert _relerr(lat.filter(s, exact_slice=True).cpu().numpy(), want) < TOL
    assert _relerr(lat.filter(s, flat=True).cpu().numpy(), want) < TOL
    rng = np.random.default_rng(3)
    add = rng.standard_normal(src.shape).astype(np.float32)
    got = lat.filter(s, sub=s, add=_cuda(add)).cpu().numpy()
    assert _relerr(got, add + (want - src)) < TOL
    got = lat.filter(s, add=_cuda(add)).cpu().numpy()
    assert _relerr(got, add + want) < TOL


@pytest.mark.parametrize("h,w,s_ij,s_rgb,C", [(150, 200, 0.1, 0.1, 96), (150, 200, 0.01, 0.125, 64)])
def test_image_features_wide(native, oracle, h, w, s_ij, s_rgb, C):
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=5), s_ij, s_rgb)
    src = np.random.default_rng(6).random((h * w, C)).astype(np.float32)
    lat = native.PermutohedralLattice(_cuda(ref))
    want = oracle.build(ref).apply(src)
    # twice on the same handle: the second call reuses the workspace and the staging choice
    for _ in range(2):
        assert _relerr(lat.filter(_cuda(src)).cpu().numpy(), want) < TOL


@pytest.mark.parametrize("n,d,C,scale", [
    (3000, 5, 64, 4.0), (2500, 5, 100, 12.0), (777, 5, 8, 2.0), (2000, 3, 48, 3.0), (1500, 7, 36, 30.0),
    (130, 5, 256, 2.0), (4100, 5, 4, 6.0),
])
def test_filter_dot_matches_composition(native, oracle, n, d, C, scale):
    """pl_lattice_apply_dot: out += coef * <filter(src * rowscale), other> against the oracle filter of the
    materialised product (the contraction the guide gradient needs, gaussian_matrix.py:450-463)."""
    src, ref = _rand(n, d, C, scale, seed=n + C)
    rng = np.random.default_rng(11)
    other = rng.standard_normal((n, C)).astype(np.float32)
    rs = rng.standard_normal((n, d)).astype(np.float32)
    out0 = rng.standard_normal((n, d)).astype(np.float32)
    lat = native.PermutohedralLattice(_cuda(ref))
    out = _cuda(out0)
    rsd = _cuda(rs)
    for j in (0, d - 1):
        lat.filter_dot(_cuda(src), _cuda(other), out[:, j], rowscale=rsd[:, j], coef=-0.5)
    lat.filter_dot(_cuda(src), _cuda(other), out[:, 0], coef=2.0)          # no row factor
    want = out0.astype(np.float64).copy()
    for j in (0, d - 1):
        f = oracle.filter(src * rs[:, j:j + 1], ref).astype(np.float64)
        want[:, j] += -0.5 * (f * other).sum(1)
    want[:, 0] += 2.0 * (oracle.filter(src, ref).astype(np.float64) * other).sum(1)
    assert _relerr(out.cpu().numpy(), want) < TOL


@pytest.mark.parametrize("n,C,scale", [(3000, 64, 4.0), (2500, 100, 12.0), (777, 8, 2.0), (130, 256, 2.0),
                                       (5000, 68, 30.0)])
def test_filter_dot_multi_matches_single(native, n, C, scale):
    """All five row factors in one pass (pl_lattice_apply_dot_multi) against five single passes."""
    src, ref = _rand(n, 5, C, scale, seed=n + C + 1)
    rng = np.random.default_rng(12)
    other = _cuda(rng.standard_normal((n, C)).astype(np.float32))
    rs = _cuda(rng.standard_normal((n, 5)).astype(np.float32))
    out0 = rng.standard_normal((n, 5)).astype(np.float32)
    lat = native.PermutohedralLattice(_cuda(ref))
    one, many = _cuda(out0), _cuda(out0)
    for j in range(5):
        lat.filter_dot(_cuda(src), other, one[:, j], rowscale=rs[:, j], coef=-1.5)
    lat.filter_dot_multi(_cuda(src), other, many, rs, coef=-1.5)
    assert _relerr(many.cpu().numpy(), one.cpu().numpy()) < 2e-5


def test_guide_gradient_contracted_path_matches_slab_path(native):
    """_grad_reference with the contracted filter (L % 4 == 0) against its own slab form (the reference's
    arithmetic per column, checked against the golden gradients elsewhere)."""
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    ref = torch.from_numpy(synthetic.ijrgb_features(synthetic.image(60, 80, seed=3))).cuda()
    g = torch.Generator(device="cuda").manual_seed(5)
    src = torch.rand((ref.shape[0], 16), generator=g, device="cuda")
    grad = torch.randn((ref.shape[0], 16), generator=g, device="cuda")
    lat = native.PermutohedralLattice(ref)
    fast, wg_fast = gm._grad_reference(lat, src, ref, grad)
    # 15 channels + a zero column: same sums, but L % 4 != 0 never takes the contracted path
    slow, wg_slow = gm._grad_reference(lat, src[:, :15].contiguous(), ref, grad[:, :15].contiguous())
    fast15, _ = gm._grad_reference(lat, torch.cat([src[:, :15], torch.zeros_like(src[:, :1])], 1).contiguous(), ref,
                                   torch.cat([grad[:, :15], torch.zeros_like(src[:, :1])], 1).contiguous())
    assert _relerr(fast15.cpu().numpy(), slow.cpu().numpy()) < 2e-4
    assert fast.shape == slow.shape and torch.isfinite(fast).all()


def test_shape_and_type_errors(native):
    ref = torch.rand(10, 3, device="cuda")
    with pytest.raises(AssertionError):
        native.lattice.filter(torch.rand(9, 2, device="cuda"), ref)
    with pytest.raises(TypeError):
        native.lattice.filter(torch.rand(10, 2, device="cuda", dtype=torch.float64), ref.double())
    with pytest.raises(Exception):
        native.PermutohedralLattice(torch.rand(10, 17, device="cuda"))


def test_full_size_properties(native):
    """BASELINE config-2 size (740x500, 64 channels): size-independent properties."""
    h, w, C = 500, 740, 64
    ref = _cuda(synthetic.ijrgb_features(synthetic.image(h, w, seed=5)))
    lat = native.PermutohedralLattice(ref)
    g = torch.Generator(device="cuda").manual_seed(0)
    a = torch.rand((h * w, C), device="cuda", generator=g)
    b = torch.rand((h * w, C), device="cuda", generator=g)
    fa, fb = lat.filter(a), lat.filter(b)
    scale = float(fa.abs().max())
    # linearity
    assert float((lat.filter(a + 2 * b) - (fa + 2 * fb)).abs().max()) < 1e-4 * scale * 3
    # channel independence: filtering a column block gives the same columns
    assert float((lat.filter(a[:, :8].contiguous()) - fa[:, :8]).abs().max()) < 1e-4 * scale
    # positivity and run-to-run agreement within the atomic-order tolerance
    assert float(fa.min()) >= 0
    assert float((lat.filter(a) - fa).abs().max()) < 1e-4 * scale
    # degree D = filter(1) > 0 (RbfLaplacian, gaussian_matrix.py:314)
    assert float(lat.filter(torch.ones((h * w, 1), device="cuda")).min()) > 0

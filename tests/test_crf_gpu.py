"""GPU parity for the operator surface: autograd functions, adjacency, mean-field CRF.

Oracle = oracle/crf_oracle.py (line-by-line torch-CPU restatement of the
reference's Python layer on the CPU oracle filter).  Tolerances: filtered
values / energies 1e-4 of output scale; argmax label identical on > 99.9 % of
pixels (north_star); gradients 5e-4 relative (the reference's own gradcheck
rtol, crf/gaussian_matrix.py:517).
"""
import importlib

import numpy as np
import pytest
import torch

synthetic = importlib.import_module("depth_estimation_b200.synthetic")

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.detach().cpu().float(), b.detach().cpu().float()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _scene(h, w, L, seed=0, s=0.1):
    img = synthetic.image(h, w, seed)
    rgb = torch.from_numpy(img).permute(2, 0, 1)[None].contiguous()          # [1,3,h,w]
    e0 = torch.from_numpy(synthetic.unary(h, w, L, seed)).reshape(h, w, L)
    logits = (-e0).permute(2, 0, 1)[None].contiguous()                      # [1,L,h,w]
    return rgb, logits


def test_softmax_compat_kernel(native):
    import ctypes as C
    N = native._native
    for n, L in [(1000, 16), (777, 64), (300, 18), (257, 128), (100, 256), (50, 7)]:
        g = torch.Generator().manual_seed(L)
        e = torch.randn((n, L), generator=g) * 3
        mu = torch.from_numpy(synthetic.charb_matrix(L)) + torch.rand((L, L), generator=g) * 0.1
        want_q = torch.softmax(-e, dim=1)
        want_p = (want_q @ mu.t()) * 1.7
        ed, mt = e.cuda(), mu.t().contiguous().cuda()
        p, q = torch.empty_like(ed), torch.empty_like(ed)
        N.check(N.lib().pl_softmax_compat(C.c_void_p(ed.data_ptr()), C.c_void_p(mt.data_ptr()),
                                          C.c_void_p(p.data_ptr()), C.c_void_p(q.data_ptr()),
                                          n, L, 1.7, N.stream_ptr(ed.device)))
        assert _rel(q, want_q) < 1e-5
        assert _rel(p, want_p) < 1e-5


@pytest.mark.parametrize("n,L", [(1000, 64), (777, 16), (333, 7), (130, 256), (65, 128), (50, 512), (40, 516), (9, 4)])
def test_softmax_backward_kernel(native, n, L):
    """pl_softmax_backward against torch autograd of scale * softmax(-E) (crf_module.py:99-102),
    in place, with the fused accumulation of the unary gradient."""
    import ctypes as C
    N = native._native
    g = torch.Generator().manual_seed(n + L)
    e = (torch.randn((n, L), generator=g) * 2).double().requires_grad_()
    dq = torch.randn((n, L), generator=g)
    q = torch.softmax(-e, dim=1)
    (q * (1.3 * dq.double())).sum().backward()
    want = e.grad
    acc0 = torch.randn((n, L), generator=g)
    dq_d, q_d, acc = dq.cuda(), q.detach().float().cuda(), acc0.cuda()
    N.check(N.lib().pl_softmax_backward(C.c_void_p(dq_d.data_ptr()), C.c_void_p(q_d.data_ptr()),
                                        C.c_void_p(dq_d.data_ptr()), C.c_void_p(acc.data_ptr()),
                                        n, L, 1.3, N.stream_ptr(dq_d.device)))
    torch.cuda.synchronize()
    assert _rel(dq_d, want) < 1e-5
    assert _rel(acc, acc0.double() + want) < 1e-5
    # out of place, no accumulator
    out = torch.empty((n, L), device="cuda")
    dq2 = dq.cuda()
    N.check(N.lib().pl_softmax_backward(C.c_void_p(dq2.data_ptr()), C.c_void_p(q_d.data_ptr()),
                                        C.c_void_p(out.data_ptr()), C.c_void_p(0), n, L, 1.3,
                                        N.stream_ptr(out.device)))
    torch.cuda.synchronize()
    assert _rel(out, want) < 1e-5
    assert torch.equal(dq2.cpu(), dq)


@pytest.mark.parametrize("n", [1, 127, 128, 1000, 40000])
def test_meanfield_backward_step_kernel(native, n):
    """pl_meanfield_backward_step (L = 64) against the composition it replaces, in fp64:
    dQ = dP Mu, dMu += dP^T Q, dE = -s Q (dQ - rowsum(dQ Q)), dE0 += dE."""
    import ctypes as C
    N = native._native
    L, s = 64, 0.7
    g = torch.Generator().manual_seed(n)
    dp = torch.randn((n, L), generator=g)
    q = torch.softmax(torch.randn((n, L), generator=g) * 2, dim=1)
    mu = torch.from_numpy(synthetic.charb_matrix(L)) + torch.rand((L, L), generator=g) * 0.1
    de0 = torch.randn((n, L), generator=g)
    dmu0 = torch.randn((L, L), generator=g)
    dq = dp.double() @ mu.double()
    want_de = -s * q.double() * (dq - (dq * q.double()).sum(1, keepdim=True))
    want_dmu = dmu0.double() + dp.double().t() @ q.double()
    dp_d, q_d, mu_d, de0_d, dmu_d = dp.cuda(), q.cuda(), mu.contiguous().cuda(), de0.cuda(), dmu0.cuda()
    N.check(N.lib().pl_meanfield_backward_step(
        C.c_void_p(dp_d.data_ptr()), C.c_void_p(q_d.data_ptr()), C.c_void_p(mu_d.data_ptr()),
        C.c_void_p(de0_d.data_ptr()), C.c_void_p(dmu_d.data_ptr()), n, L, s, N.stream_ptr(dp_d.device)))
    torch.cuda.synchronize()
    assert _rel(dp_d, want_de) < 1e-5
    assert _rel(de0_d, de0.double() + want_de) < 1e-5
    assert _rel(dmu_d, want_dmu) < 1e-5
    assert N.lib().pl_meanfield_backward_step(
        C.c_void_p(dp_d.data_ptr()), C.c_void_p(q_d.data_ptr()), C.c_void_p(mu_d.data_ptr()),
        C.c_void_p(0), C.c_void_p(0), n, 32, s, N.stream_ptr(dp_d.device)) != 0     # L = 64 only


@pytest.mark.parametrize("L", [16, 32, 64])
def test_softmax_compat_tensor_core_path(native, L):
    """tcgen05 3xTF32 compatibility transform (pl_fused_tc.cu, MODE 2) vs fp64 and vs the
    CUDA-core kernel: full and ragged tiles, with and without the Q output."""
    import ctypes as C
    N = native._native
    for n in (1, 127, 128, 129, 1000, 4099):
        g = torch.Generator().manual_seed(L + n)
        e = torch.randn((n, L), generator=g) * 3
        mu = torch.from_numpy(synthetic.charb_matrix(L)) + torch.rand((L, L), generator=g) * 0.1
        want_q = torch.softmax(-e.double(), dim=1)
        want_p = (want_q @ mu.double().t()) * 1.7
        ed, mt = e.cuda(), mu.t().contiguous().cuda()
        outs = {}
        for name, flags, with_q in (("tc", 0, True), ("tc_noq", 0, False),
                                    ("cuda", N.PL_MEANFIELD_NO_TENSOR, True)):
            p, q = torch.full_like(ed, float("nan")), torch.full_like(ed, float("nan"))
            N.check(N.lib().pl_softmax_compat_ex(
                C.c_void_p(ed.data_ptr()), C.c_void_p(mt.data_ptr()), C.c_void_p(p.data_ptr()),
                C.c_void_p(q.data_ptr() if with_q else 0), n, L, 1.7, flags, N.stream_ptr(ed.device)))
            torch.cuda.synchronize()
            outs[name] = p
            assert _rel(p.double(), want_p) < 2e-6, (name, n, L)
            if with_q:
                assert _rel(q.double(), want_q) < 2e-6
        assert N.lib().pl_debug_tc_timeouts() == 0
        # the split-TF32 result is as close to fp64 as the fp32 FMA chain is
        assert _rel(outs["tc"], outs["cuda"]) < 4e-6


def test_layout_helpers(native):
    import ctypes as C
    N = native._native
    x = torch.randn((2, 5, 70), device="cuda")
    y = torch.empty((2, 70, 5), device="cuda")
    N.check(N.lib().pl_nchw_to_nlc(C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), 2, 5, 70,
                                   -1.0, N.stream_ptr(x.device)))
    assert torch.equal(y, -x.permute(0, 2, 1))
    z = torch.empty_like(x)
    N.check(N.lib().pl_nlc_to_nchw(C.c_void_p(y.data_ptr()), C.c_void_p(z.data_ptr()), 2, 5, 70,
                                   -1.0, N.stream_ptr(x.device)))
    assert torch.equal(z, x)


def test_lattice_filter_autograd_matches_oracle(native):
    """Forward, grad wrt source, and grad wrt reference (the 2L(1+d)-channel path)."""
    from oracle import crf_oracle as O
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    g = torch.Generator().manual_seed(0)
    n, L, d = 400, 3, 3
    src = torch.rand((n, L), generator=g)
    ref = torch.rand((n, d), generator=g) * 3
    gout = torch.randn((n, L), generator=g)

    s0, r0 = src.clone().requires_grad_(), ref.clone().requires_grad_()
    out0 = O.LatticeFilter.apply(s0, r0)
    out0.backward(gout)
    s1, r1 = src.cuda().requires_grad_(), ref.cuda().requires_grad_()
    out1 = gm.LatticeFilter.apply(s1, r1)
    out1.backward(gout.cuda())
    assert _rel(out1, out0) < 1e-4
    assert _rel(s1.grad, s0.grad) < 5e-4
    assert _rel(r1.grad, r0.grad) < 5e-4

    # source-only branch (gaussian_matrix.py:445-446)
    s2 = src.cuda().requires_grad_()
    gm.LatticeFilter.apply(s2, ref.cuda()).backward(gout.cuda())
    want = O.latticefilter(gout, ref)
    assert _rel(s2.grad, want) < 1e-4


def test_adjacency_and_laplacians(native, oracle):
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    rgb, logits = _scene(40, 56, 6, seed=3)
    from oracle import crf_oracle as O
    guide = O.ijrgb_guide(rgb)
    src = torch.softmax(logits, 1)
    want = O.batched_adjacency(src, guide)
    got = gm.BatchedAdjacency()(src.cuda(), guide.cuda())
    assert _rel(got, want) < 1e-4
    ref = guide.reshape(5, -1).t().contiguous()
    U = torch.rand((ref.shape[0], 2))
    D = torch.from_numpy(oracle.filter(np.ones((ref.shape[0], 1), np.float32), ref.numpy()))
    lap = gm.RbfLaplacian(ref.cuda(), normalize=True)
    want = U - torch.from_numpy(oracle.filter((U / D.sqrt()).numpy(), ref.numpy())) / D.sqrt()
    assert _rel(lap @ U.cuda(), want) < 2e-4
    W = gm.LatticeGaussian(ref.cuda())
    assert _rel(W @ U.cuda(), torch.from_numpy(oracle.filter(U.numpy(), ref.numpy())) - U) < 1e-4


@pytest.mark.parametrize("h,w,L,niters", [(48, 64, 16, 5), (60, 80, 64, 3), (32, 40, 18, 2)])
def test_crf_forward_fused_and_composed_match_oracle(native, h, w, L, niters):
    from oracle import crf_oracle as O
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    rgb, logits = _scene(h, w, L, seed=L)
    guide = O.ijrgb_guide(rgb)
    gamma, s = torch.tensor(0.05), torch.tensor(0.3)
    want = O.crf_as_rnn(guide, logits, gamma, s, niters)

    crf = cm.CRFasRNN(cm.charb(0.05), niters=niters).cuda()
    with torch.no_grad():
        crf.Mu.s.fill_(0.3)
        got_fused = crf(guide.cuda(), logits.cuda())
    crf.fused = False
    with torch.no_grad():
        got_comp = crf(guide.cuda(), logits.cuda())
    for got in (got_fused, got_comp):
        assert _rel(got, want) < 2e-4
        agree = (got.cpu().argmax(1) == want.argmax(1)).float().mean()
        assert float(agree) > 0.999


def test_crf_backward_matches_oracle(native):
    """Trainable-CRF backward (config 2 shape in miniature): grads wrt logits, gamma, s, guide scales."""
    from oracle import crf_oracle as O
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    h, w, L, niters = 24, 32, 8, 2
    rgb, logits = _scene(h, w, L, seed=11)
    target = torch.from_numpy(synthetic.disparity(h, w, L, 11)).float()[None, None]

    # oracle side
    lg0 = logits.clone().requires_grad_()
    gamma0, s0 = torch.tensor(0.05, requires_grad=True), torch.tensor(0.0, requires_grad=True)
    sij0, srgb0 = torch.tensor(0.1, requires_grad=True), torch.tensor(0.1, requires_grad=True)
    ij = O.ijrgb_guide(rgb, 1.0, 1.0)[:, :2]
    guide0 = torch.cat([ij / sij0, rgb / srgb0], dim=1)
    out0 = O.crf_as_rnn(guide0, lg0, gamma0, s0, niters)
    loss0 = (O.logits2average_depth(out0) - target).abs().mean()
    loss0.backward()

    # product side
    crf = cm.CRFasRNN(cm.charb(0.05), niters=niters).cuda()
    guide_mod = cm.ijrgbGuide(0.1, 0.1).cuda()
    mb = importlib.import_module("depth_estimation_b200.crf.mb_stereo_crf")
    lg1 = logits.cuda().requires_grad_()
    out1 = crf(guide_mod(rgb.cuda()), lg1)
    loss1 = (mb.logits2average_depth(out1) - target.cuda()).abs().mean()
    loss1.backward()

    assert abs(float(loss1) - float(loss0)) < 1e-4 * max(1.0, abs(float(loss0)))
    assert _rel(lg1.grad, lg0.grad) < 2e-3
    for got, want in [(crf.Mu.gamma.grad, gamma0.grad), (crf.Mu.s.grad, s0.grad),
                      (guide_mod.s_ij.grad, sij0.grad), (guide_mod.s_rgb.grad, srgb0.grad)]:
        assert abs(float(got) - float(want)) <= 5e-3 * max(abs(float(want)), 1e-6) + 1e-7


def test_stereo_wrappers_run(native):
    mb = importlib.import_module("depth_estimation_b200.crf.mb_stereo_crf")
    torch.manual_seed(0)
    up = mb.CRFdepthUpsampler(niters=2).cuda()
    disp = torch.rand((1, 1, 12, 16), device="cuda") * 10
    img = torch.rand((1, 3, 48, 64), device="cuda")
    with torch.no_grad():
        out = up((disp, img, None))
    assert out.shape == (1, 1, 48, 64) and torch.isfinite(out).all()
    refiner = mb.CRFdepthRefiner(d_in=8, d_guide=5, niters=2).cuda()
    with torch.no_grad():
        out = refiner((torch.randn((1, 6, 24, 32), device="cuda"), torch.rand((1, 3, 24, 32), device="cuda"),
                       torch.randn((1, 8, 24, 32), device="cuda")))
    assert out.shape == (1, 1, 24, 32) and torch.isfinite(out).all()


@pytest.mark.parametrize("h,w,L,niters,s_ij", [(37, 53, 4, 1, 0.1), (48, 64, 16, 5, 0.1), (50, 70, 40, 2, 0.1),
                                                 (64, 96, 64, 3, 0.1), (40, 60, 64, 4, 0.01),
                                                 (150, 201, 64, 3, 0.1), (240, 321, 32, 3, 0.1),
                                                 (90, 130, 32, 2, 0.02)])
def test_fused_iteration_matches_unfused_and_oracle(native, oracle, h, w, L, niters, s_ij):
    """pl_meanfield: fused per-iteration kernel (pl_fused.cu) vs separate kernels vs CPU oracle."""
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=L), s_ij, 0.1)
    e0 = synthetic.unary(h, w, L, seed=L)
    mu = synthetic.charb_matrix(L) * 0.7
    # oracle, pixel-major (crf_module.py:97-103 with W = filter - id)
    hdl = oracle.build(ref)
    e = e0.copy()
    for _ in range(niters):
        x = -e - (-e).max(1, keepdims=True)
        q = np.exp(x) / np.exp(x).sum(1, keepdims=True)
        p = (q @ mu.T).astype(np.float32) * np.float32(1.3)
        e = e0 + (hdl.apply(p) - p)
    lat = native.PermutohedralLattice(torch.from_numpy(ref).cuda())
    e0d, mud = torch.from_numpy(e0).cuda(), torch.from_numpy(mu).cuda()
    fused = lat.mean_field(e0d, mud, niters, 1.3).cpu().numpy()
    unfused = lat.mean_field(e0d, mud, niters, 1.3, unfused=True).cpu().numpy()
    flat = lat.mean_field(e0d, mud, niters, 1.3, flat=True).cpu().numpy()
    cuda_core = lat.mean_field(e0d, mud, niters, 1.3, no_tensor=True).cpu().numpy()
    tile = lat.mean_field(e0d, mud, niters, 1.3, tile_kernel=True).cpu().numpy()
    assert native._native.lib().pl_debug_tc_timeouts() == 0
    scale = np.abs(e).max()
    for got in (fused, unfused, flat, cuda_core, tile):
        assert np.abs(got - e).max() / scale < 2e-4
        assert (got.argmin(1) == e.argmin(1)).mean() > 0.999


def test_guide_gradient_survives_an_earlier_no_grad_call_on_the_same_tensor(native):
    """The lattice cache holds only detached values: an inference call on a guide tensor must not
    make a later training call on the SAME tensor lose its guide gradient (the cache used
    to hand back the grad-less flattened copy), for the fused autograd op and for the composed path
    through BatchedAdjacency / LatticeFilter alike."""
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    rgb, logits = _scene(24, 32, 8, seed=3)
    guide = cm.ijrgbGuide(0.1, 0.1, trainable=False)(rgb.cuda()).clone().requires_grad_()   # a leaf that wants grad
    lg = logits.cuda()
    for fused in (True, False):
        gm.clear_lattice_cache()
        crf = cm.CRFasRNN(cm.charb(0.05), niters=2, fused=fused).cuda()
        with torch.no_grad():
            ref_out = crf(guide, lg)                      # fills the cache without autograd
        guide.grad = None
        out = crf(guide, lg)
        assert _rel(out, ref_out) < 2e-4
        out.square().mean().backward()
        assert guide.grad is not None and torch.isfinite(guide.grad).all()
        assert float(guide.grad.abs().max()) > 0
        g_cached = guide.grad.clone()
        # same call with a cold cache gives the same gradient
        gm.clear_lattice_cache()
        guide.grad = None
        crf(guide, lg).square().mean().backward()
        assert _rel(g_cached, guide.grad) < 1e-3


def test_scalar_confidence_is_applied_with_autograd_on(native):
    """CRFasRNN.depth with a python-scalar confidence: the unfused (autograd) branch used to drop it."""
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    rgb, logits = _scene(24, 32, 8, seed=4)
    guide = cm.ijrgbGuide(0.1, 0.1, trainable=False)(rgb.cuda())
    crf = cm.CRFasRNN(cm.charb(0.05), niters=2).cuda()
    with torch.no_grad():
        want = crf.depth(guide, logits.cuda(), confidence=0.5)
    lg = logits.cuda().requires_grad_()
    got = crf.depth(guide, lg, confidence=0.5)
    assert _rel(got, want) < 1e-3
    assert _rel(got, crf.depth(guide, lg)) > 1e-3          # and it is not the confidence = 1 result

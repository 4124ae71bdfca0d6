"""The rows either side of the mean-field loop on the GPU:
guide features fused into the lattice build, the plane-sweep cost volume, and the depth
read-out fused into the last slice.  Checked against the golden vectors of the real reference
modules, the CPU oracles, and the unfused CUDA path.
"""
import importlib
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
synthetic = importlib.import_module("depth_estimation_b200.synthetic")


def _rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


# ---- f1: cost volume (crf/depth.py:34-53) ---------------------------------------------------

def test_cost_volume_vs_reference_golden(native):
    D = importlib.import_module("depth_estimation_b200.crf.depth")
    z = np.load(os.path.join(G, "cost_volume_26x44.npz"))
    left, right = z["left"], z["right"]
    # numpy in -> numpy out, the reference's call shape; fp32 arithmetic vs the reference's fp64
    assert _rel(D.disparity_badness(left, right), z["ad9"]) < 1e-5
    assert _rel(D.disparity_badness(left, right, 5, D.SD), z["sd5"]) < 1e-5
    assert _rel(D.disparity_badness(left, right, 3, D.nprod), z["np3"]) < 1e-5
    est = D.disparity_estimate(left, right)
    assert est.shape == z["est"].shape and (est == z["est"]).mean() > 0.999
    # CUDA tensors in -> CUDA tensor out, pixel-major view usable as E0
    vol = D.disparity_badness(torch.from_numpy(left).float().cuda(), torch.from_numpy(right).float().cuda())
    assert vol.is_cuda and vol.shape == z["ad9"].shape and vol.is_contiguous()
    assert _rel(vol.cpu(), z["ad9"]) < 1e-5


@pytest.mark.parametrize("h,w,c,L,ws,crit", [
    (5, 70, 3, 11, 9, "AD"),        # image lower than the window: double reflection rows
    (67, 130, 3, 21, 9, "AD"),      # ragged tile edges, L not a multiple of 8
    (40, 64, 1, 8, 3, "SD"),        # single channel, exact tile width
    (33, 97, 3, 16, 17, "nprod"),   # largest window
    (200, 333, 3, 55, 9, "AD"),
])
def test_cost_volume_vs_oracle(native, h, w, c, L, ws, crit):
    D = importlib.import_module("depth_estimation_b200.crf.depth")
    from oracle import depth_oracle as O
    rng = np.random.default_rng(h * w + L)
    a, b = rng.random((h, w, c)), rng.random((h, w, c))
    want = O.disparity_badness(a, b, ws, getattr(O, crit), max_disp=L)
    got = D.cost_volume(torch.from_numpy(a).float().cuda(), torch.from_numpy(b).float().cuda(), L, ws,
                        getattr(D, crit))
    assert got.shape == (h, w, L)
    assert _rel(got.cpu(), want) < 1e-5
    with pytest.raises(RuntimeError):
        D.cost_volume(torch.zeros((4, 4, 3), device="cuda"), torch.zeros((4, 4, 3), device="cuda"), 2, 4)


# ---- a1: ijrgbGuide evaluated inside the elevation kernel (crf_module.py:106-114) -----------

@pytest.mark.parametrize("h,w,s_ij,s_rgb", [(20, 28, 0.1, 0.1), (61, 47, 0.05, 0.125), (128, 190, 0.1, 0.07)])
def test_lattice_from_rgb_is_bit_identical_to_the_guide_module(native, h, w, s_ij, s_rgb):
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    rgb = torch.from_numpy(synthetic.image(h, w, seed=h)).permute(2, 0, 1).contiguous().cuda()
    guide = cm.ijrgbGuide(s_ij, s_rgb).cuda()
    with torch.no_grad():
        feats = guide(rgb[None])[0].reshape(5, -1).t().contiguous()      # [n, 5] as the module lays it
    a = native.PermutohedralLattice(feats)
    b = native.PermutohedralLattice.from_rgb(rgb, s_ij, s_rgb)
    assert a.num_vertices == b.num_vertices
    ea, eb = a.export(), b.export()
    for k in ("greedy", "rank", "bary", "vid", "keys", "nbr"):
        assert torch.equal(ea[k], eb[k]), k
    src = torch.rand((h * w, 8), device="cuda")
    # identical lattices; the splat's atomic adds are not ordered, so values agree to rounding
    assert _rel(a.filter(src).cpu(), b.filter(src).cpu()) < 1e-6


def test_lattice_from_rgb_matches_reference_golden_guide(native):
    """The guide tensor in the golden file was produced by the reference's ijrgbGuide."""
    z = np.load(os.path.join(G, "crf_as_rnn_20x28.npz"))
    rgb = torch.from_numpy(z["rgb"])[0].cuda()
    feats = torch.from_numpy(z["guide"])[0].reshape(5, -1).t().contiguous().cuda()
    a = native.PermutohedralLattice(feats)
    b = native.PermutohedralLattice.from_rgb(rgb, 0.1, 0.1)
    ea, eb = a.export(), b.export()
    for k in ("greedy", "rank", "bary", "vid", "keys", "nbr"):
        assert torch.equal(ea[k], eb[k]), k


# ---- f2: depth read-out fused into the last slice (mb_stereo_crf.py:62-66) ------------------

@pytest.mark.parametrize("h,w,L,niters,labels", [(48, 64, 16, 3, False), (64, 96, 64, 5, False),
                                                 (50, 70, 40, 2, True), (37, 53, 5, 2, False),
                                                 (60, 80, 128, 2, True), (30, 40, 64, 0, False)])
def test_fused_depth_readout_matches_unfused(native, h, w, L, niters, labels):
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=L), 0.1, 0.1)
    e0 = torch.from_numpy(synthetic.unary(h, w, L, seed=L)).cuda()
    mu = torch.from_numpy(synthetic.charb_matrix(L) * 0.7).cuda()
    lab = torch.linspace(0.5, 9.0, L, device="cuda") if labels else None
    lat = native.PermutohedralLattice(torch.from_numpy(ref).cuda())
    e = lat.mean_field(e0, mu, niters, 1.3)
    q = torch.softmax(-e.double(), dim=1)
    lv = lab.double() if labels else torch.arange(L, device="cuda", dtype=torch.float64)
    want = (q * lv[None]).sum(1)
    depth, label = lat.mean_field_depth(e0, mu, niters, 1.3, labels=lab, want_label=True)
    assert depth.shape == (h * w,) and label.dtype == torch.int32
    # two runs of the loop differ by the order of the splat's atomic adds (~1e-7 per iteration,
    # amplified by the soft-argmax at near-tie pixels): 3e-4 of the depth range across runs,
    # 2e-5 for the read-out of one and the same E below
    assert _rel(depth.cpu(), want.cpu()) < 3e-4
    assert (label.long() == e.argmin(1)).float().mean() > 0.999
    only_label = lat.mean_field_depth(e0, mu, niters, 1.3, want_depth=False, want_label=True)
    assert only_label[0] is None and torch.equal(only_label[1], label)
    # the read-out of rows already in memory: energies (+1) and logits (-1)
    import ctypes as C
    N = native._native
    d2 = torch.empty_like(depth)
    N.check(N.lib().pl_softargmax(C.c_void_p(e.data_ptr()), h * w, L,
                                  C.c_void_p(lab.data_ptr() if labels else 0), 1.0,
                                  C.c_void_p(d2.data_ptr()), C.c_void_p(0), N.stream_ptr(e.device)))
    assert _rel(d2.cpu(), want.cpu()) < 2e-5
    neg = (-e).contiguous()
    N.check(N.lib().pl_softargmax(C.c_void_p(neg.data_ptr()), h * w, L,
                                  C.c_void_p(lab.data_ptr() if labels else 0), -1.0,
                                  C.c_void_p(d2.data_ptr()), C.c_void_p(0), N.stream_ptr(e.device)))
    assert _rel(d2.cpu(), want.cpu()) < 2e-5


def test_stereo_pair_to_depth_matches_reference_composition(native):
    """images -> cost volume -> CRF -> depth, every stage on the GPU, against the same pipeline
    assembled from the CPU oracles (crf/depth.py + crf_module.py:93-104 + mb_stereo_crf.py:62-66)."""
    from oracle import depth_oracle as DO
    from oracle.lattice_oracle import Oracle
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    h, w, niters = 40, 96, 3
    L = w // 6
    left = synthetic.image(h, w, seed=11).astype(np.float64)
    right = np.roll(left, -3, axis=1)
    e0 = DO.disparity_badness(left, right).reshape(h * w, L).astype(np.float32)
    feats = synthetic.ijrgb_features(left.astype(np.float32), 0.1, 0.1)
    mu = synthetic.charb_matrix(L, 0.05)
    hdl = Oracle(faithful=False).build(feats)
    e = e0.copy()
    for _ in range(niters):
        x = -e - (-e).max(1, keepdims=True)
        qq = np.exp(x) / np.exp(x).sum(1, keepdims=True)
        p = (qq @ mu.T).astype(np.float32)
        e = e0 + (hdl.apply(p) - p)
    x = -e.astype(np.float64)
    qq = np.exp(x - x.max(1, keepdims=True))
    want = (qq / qq.sum(1, keepdims=True) * np.arange(L)[None]).sum(1)

    got = cm.stereo_pair_depth(torch.from_numpy(left).float().cuda(), torch.from_numpy(right).float().cuda(),
                               niters=niters, gamma=0.05, s_ij=0.1, s_rgb=0.1)
    assert got.shape == (h, w)
    assert np.abs(got.cpu().numpy().reshape(-1) - want).max() < 2e-3 * L
    assert np.abs(got.cpu().numpy().reshape(-1) - want).mean() < 1e-4 * L


def test_module_depth_matches_logits2average_depth(native):
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    mb = importlib.import_module("depth_estimation_b200.crf.mb_stereo_crf")
    torch.manual_seed(3)
    B, L, H, W = 2, 16, 24, 40
    rgb = torch.rand((B, 3, H, W), device="cuda")
    logits = torch.randn((B, L, H, W), device="cuda") * 2
    conf = torch.rand((B, 1, H, W), device="cuda") * 0.5 + 0.75
    guide = cm.ijrgbGuide(0.1, 0.1).cuda()
    crf = cm.CRFasRNN(cm.charb(0.05), niters=3).cuda()
    labels = torch.linspace(0, 4, L, device="cuda")
    with torch.no_grad():
        g = guide(rgb)
        want = mb.logits2average_depth(crf(g, logits, conf))
        got = crf.depth(g, logits, conf)
        assert got.shape == (B, 1, H, W) and _rel(got.cpu(), want.cpu()) < 2e-5
        want_l = mb.logits2average_depth(crf(g, logits, conf, labels), labels[None, :, None, None])
        assert _rel(crf.depth(g, logits, conf, labels).cpu(), want_l.cpu()) < 2e-5
    # with autograd the same call runs the differentiable composition
    lg = logits.clone().requires_grad_()
    d = crf.depth(guide(rgb), lg, conf)
    d.sum().backward()
    assert lg.grad is not None and torch.isfinite(lg.grad).all()
    assert _rel(d.detach().cpu(), want.cpu()) < 2e-4

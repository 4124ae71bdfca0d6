"""The CUDA path against the golden vectors produced by the REAL reference modules
(tests/golden/make_golden.py).  Integer artefacts bit-exact; floats within the tolerances
north_star states (1e-4 of the output scale for filtered values / energies, identical argmax
on > 99.9 % of pixels; gradients 5e-4 like the reference's own gradcheck rtol).
"""
import importlib
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(G, name))


def _rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def test_integer_artefacts_bit_exact_vs_reference_golden(native):
    z = _load("lattice_state_24x32.npz")
    lat = native.PermutohedralLattice(torch.from_numpy(z["ref"]).cuda())
    assert lat.num_vertices == int(z["M"])
    ex = {k: v.cpu().numpy() for k, v in lat.export().items()}
    for k in ("greedy", "rank", "vid", "keys", "nbr"):
        assert np.array_equal(ex[k], z[k]), k
    assert np.array_equal(ex["bary"], z["bary"])
    got = lat.filter(torch.from_numpy(z["src"]).cuda()).cpu().numpy()
    assert _rel(got, z["filtered"]) < 1e-4
    one_shot = native.lattice.filter(torch.from_numpy(z["src"]).cuda(), torch.from_numpy(z["ref"]).cuda())
    assert _rel(one_shot.cpu().numpy(), z["filtered"]) < 1e-4
    # the reference's own call shape: CPU tensors in, CPU tensor out
    host = native.lattice.filter(torch.from_numpy(z["src"]), torch.from_numpy(z["ref"]))
    assert not host.is_cuda and _rel(host.numpy(), z["filtered"]) < 1e-4


def test_lattice_filter_autograd_vs_reference_golden(native):
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    z = _load("lattice_filter_autograd.npz")
    s = torch.from_numpy(z["src"]).cuda().requires_grad_()
    r = torch.from_numpy(z["ref"]).cuda().requires_grad_()
    y = gm.LatticeFilter.apply(s, r)
    y.backward(torch.from_numpy(z["gout"]).cuda())
    assert _rel(y.detach().cpu(), z["out"]) < 1e-4
    assert _rel(s.grad.cpu(), z["grad_src"]) < 5e-4
    assert _rel(r.grad.cpu(), z["grad_ref"]) < 5e-4


def test_lattice_operators_vs_reference_golden(native):
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    z = _load("lattice_operators.npz")
    ref, U = torch.from_numpy(z["ref"]).cuda(), torch.from_numpy(z["U"]).cuda()
    got = {"gaussian": gm.LatticeGaussian(ref) @ U,
           "lap_norm": gm.RbfLaplacian(ref, normalize=True) @ U,
           "lap_unnorm": gm.RbfLaplacian(ref, normalize=False) @ U,
           "lapc_sym": gm.RbfLaplacianC(ref, 'sym') @ U,
           "lapc_right": gm.RbfLaplacianC(ref, 'right') @ U,
           "lapc_none": gm.RbfLaplacianC(ref, None) @ U}
    for k, v in got.items():
        assert _rel(v.cpu(), z[k]) < 2e-4, k


def test_crf_as_rnn_forward_backward_vs_reference_golden(native):
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    z = _load("crf_as_rnn_20x28.npz")
    L, niters = z["logits"].shape[1], int(z["niters"])
    rgb = torch.from_numpy(z["rgb"]).cuda()
    conf = torch.from_numpy(z["confidence"]).cuda()
    guide_mod = cm.ijrgbGuide(0.1, 0.1).cuda()
    crf = cm.CRFasRNN(cm.charb(float(z["gamma"])), niters=niters).cuda()
    with torch.no_grad():
        guide = guide_mod(rgb)
        assert _rel(guide.cpu(), z["guide"]) < 1e-6
        fused = crf(guide, torch.from_numpy(z["logits"]).cuda(), confidence=conf)
    assert _rel(fused.cpu(), z["out_logits"]) < 2e-4
    agree = (fused.cpu().numpy().argmax(1) == z["out_logits"].argmax(1)).mean()
    assert agree > 0.999
    # trainable path: gradients wrt logits, gamma, s and the guide scales
    lg = torch.from_numpy(z["logits"]).cuda().requires_grad_()
    out = crf(guide_mod(rgb), lg, confidence=conf)
    assert _rel(out.detach().cpu(), z["out_logits"]) < 2e-4
    labels = torch.arange(L, dtype=torch.float32, device="cuda")
    depth = (torch.softmax(out, dim=1) * labels[None, :, None, None]).sum(1)[:, None]
    loss = (depth - torch.from_numpy(z["target"]).cuda()).abs().mean()
    loss.backward()
    assert abs(float(loss) - float(z["loss"])) < 1e-4 * max(1.0, abs(float(z["loss"])))
    assert _rel(lg.grad.cpu(), z["grad_logits"]) < 2e-3
    for got, want in [(crf.Mu.gamma.grad, z["grad_gamma"]), (crf.Mu.s.grad, z["grad_s"]),
                      (guide_mod.s_ij.grad, z["grad_s_ij"]), (guide_mod.s_rgb.grad, z["grad_s_rgb"])]:
        assert abs(float(got) - float(want)) <= 5e-3 * max(abs(float(want)), 1e-6) + 1e-7


def test_matrix_form_and_small_pieces_vs_reference_golden(native):
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    mb = importlib.import_module("depth_estimation_b200.crf.mb_stereo_crf")
    z = _load("crf_matrix_form.npz")
    ref = torch.from_numpy(z["ref"]).cuda()
    lab = torch.from_numpy(z["labels"]).cuda()
    Mu = cm.compatibility_matrix(cm.charbonneir, lab)
    assert _rel(Mu.cpu(), z["Mu"]) < 1e-6
    assert _rel(cm.compatibility_matrix(cm.charbonneir2, lab).cpu(), z["Mu2"]) < 1e-6
    Q = cm.mean_field_infer(torch.from_numpy(z["E0"]).cuda(), gm.LatticeGaussian(ref), Mu, niters=4)
    assert _rel(Q.cpu(), z["Q"]) < 1e-4
    x = torch.from_numpy(z["potts_x"]).cuda()
    assert _rel(cm.potts(5).cuda()(x).detach().cpu(), z["potts_y"]) < 1e-6
    en = cm.charb(0.05).cuda().get_energies_from_scalar(torch.from_numpy(z["disp"]).cuda(),
                                                        lab[None, :, None, None])
    assert _rel(en.detach().cpu(), z["energies"]) < 1e-6
    assert _rel(cm.ijGuide(0.2).cuda()(x).detach().cpu(), z["ij_guide"]) < 1e-6
    w = _load("stereo_wrappers.npz")
    lg = torch.from_numpy(w["logits"]).cuda()
    assert _rel(mb.logits2average_depth(lg).cpu(), w["depth"]) < 1e-6
    labels = torch.linspace(0, 2, 5, device="cuda")[None, :, None, None]
    assert _rel(mb.logits2average_depth(lg, labels).cpu(), w["depth_labels"]) < 1e-6


def _load_state(module, z, prefix):
    sd = {k[len(prefix):].replace("__", "."): torch.from_numpy(np.asarray(z[k])) for k in z.files
          if k.startswith(prefix) and "_out" not in k}
    missing, unexpected = module.load_state_dict(sd, strict=True)
    assert not missing and not unexpected


@pytest.mark.parametrize("grad", [False, True])
def test_stereo_modules_vs_reference_golden(native, grad):
    """CRFdepthRefiner (16-channel guide -> d = 16 lattice), CRFwUncertainty (confidence map) and
    CRFdepthUpsampler (18 linspace labels) forwards against the outputs of the reference's own
    classes (mb_stereo_crf.py:68-102,138-163; tests/golden/make_golden.py stereo_modules), through
    the fused no-grad path and through the autograd path."""
    mb = importlib.import_module("depth_estimation_b200.crf.mb_stereo_crf")
    z = _load("stereo_wrappers.npz")
    t = lambda k: torch.from_numpy(z[k]).cuda()
    ctx = torch.enable_grad() if grad else torch.no_grad()
    tol = 5e-4      # of the depth range (soft-argmax over energies that are within 2e-4)
    with ctx:
        m = mb.CRFdepthRefiner(d_in=8, d_guide=16, niters=2, gamma=.05)
        _load_state(m, z, "refiner__")
        m = m.cuda()
        out = m((t("refiner_logits"), t("refiner_img"), t("refiner_feats")))
        assert out.shape == z["refiner_out"].shape
        assert _rel(out.detach().cpu(), z["refiner_out"]) < tol

        m = mb.CRFwUncertainty(d_in=8, d_guide=16, niters=2, gamma=.05)
        _load_state(m, z, "uncert__")
        m = m.cuda()
        torch.backends.cudnn.allow_tf32 = False
        out, conf = m((t("refiner_logits"), t("refiner_img"), t("refiner_feats")))
        assert _rel(conf.detach().cpu(), z["uncert_conf"]) < 1e-4
        assert _rel(out.detach().cpu(), z["uncert_out"]) < tol

        m = mb.CRFdepthUpsampler(niters=2, gamma=.05)
        _load_state(m, z, "upsampler__")
        m = m.cuda()
        out = m((t("upsampler_disp"), t("upsampler_img"), None))
        assert _rel(out.detach().cpu(), z["upsampler_out"]) < tol
        # second frame, different max_depth, same module: the labels (and Mu) must be rebuilt
        out2 = m((t("upsampler_disp") * 1.7, t("upsampler_img"), None))
        assert _rel(out2.detach().cpu(), z["upsampler_out2"]) < tol

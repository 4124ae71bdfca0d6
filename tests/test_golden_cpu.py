"""The CPU oracle against the golden vectors produced by the REAL reference modules
(tests/golden/make_golden.py: crf/gaussian_matrix.py, crf/crf_module.py, crf/mb_stereo_crf.py
executed through import stubs on the unmodified permutohedral.h).  This is what pins
oracle/lattice_oracle.c and oracle/crf_oracle.py; no GPU involved.
"""
import os

import numpy as np
import pytest
import torch

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    path = os.path.join(G, name)
    if not os.path.exists(path):
        pytest.skip(f"{name} missing (run tests/golden/make_golden.py where /root/reference exists)")
    return np.load(path)


def test_oracle_integer_state_matches_reference_golden(oracle):
    z = _load("lattice_state_24x32.npz")
    st = oracle.state(z["ref"])
    assert st.M == int(z["M"])
    for k in ("greedy", "rank", "vid", "keys", "nbr"):
        assert np.array_equal(getattr(st, k), z[k]), k
    assert np.array_equal(st.bary, z["bary"])                       # bit-exact floats too
    assert np.array_equal(oracle.filter(z["src"], z["ref"]), z["filtered"])


def test_oracle_autograd_matches_reference_golden():
    from oracle import crf_oracle as O
    z = _load("lattice_filter_autograd.npz")
    s = torch.from_numpy(z["src"]).requires_grad_()
    r = torch.from_numpy(z["ref"]).requires_grad_()
    y = O.LatticeFilter.apply(s, r)
    y.backward(torch.from_numpy(z["gout"]))
    assert np.array_equal(y.detach().numpy(), z["out"])
    np.testing.assert_allclose(s.grad.numpy(), z["grad_src"], rtol=0, atol=0)
    np.testing.assert_allclose(r.grad.numpy(), z["grad_ref"], rtol=1e-6, atol=1e-6)


def test_oracle_crf_matches_reference_golden():
    from oracle import crf_oracle as O
    z = _load("crf_as_rnn_20x28.npz")
    rgb = torch.from_numpy(z["rgb"])
    guide = O.ijrgb_guide(rgb)
    assert np.array_equal(guide.numpy(), z["guide"])
    L = z["logits"].shape[1]
    out = O.crf_as_rnn(guide, torch.from_numpy(z["logits"]), torch.tensor(float(z["gamma"])),
                       torch.tensor(0.0), int(z["niters"]), confidence=torch.from_numpy(z["confidence"]),
                       labels=torch.arange(L).float())
    np.testing.assert_allclose(out.numpy(), z["out_logits"], rtol=0, atol=1e-5)
    depth = O.logits2average_depth(out)
    np.testing.assert_allclose(depth.numpy(), z["depth"], rtol=0, atol=1e-5)


def test_oracle_matrix_form_matches_reference_golden():
    from oracle import crf_oracle as O
    z = _load("crf_matrix_form.npz")
    Q = O.mean_field_infer(torch.from_numpy(z["E0"]), torch.from_numpy(z["ref"]),
                           torch.from_numpy(z["Mu"]), niters=4)
    np.testing.assert_allclose(Q.numpy(), z["Q"], rtol=0, atol=1e-6)
    lab = torch.from_numpy(z["labels"])
    np.testing.assert_allclose(O.charbonneir(lab[:, None], lab[None, :]).numpy(), z["Mu"], atol=1e-7)


def test_oracle_stereo_readout_matches_reference_golden():
    from oracle import crf_oracle as O
    z = _load("stereo_wrappers.npz")
    lg = torch.from_numpy(z["logits"])
    np.testing.assert_allclose(O.logits2average_depth(lg).numpy(), z["depth"], atol=1e-6)
    lab = torch.linspace(0, 2, 5)[None, :, None, None]
    np.testing.assert_allclose(O.logits2average_depth(lg, lab).numpy(), z["depth_labels"], atol=1e-6)


def test_oracle_stereo_modules_match_reference_golden():
    """CRFdepthRefiner / CRFwUncertainty / CRFdepthUpsampler forwards (mb_stereo_crf.py:68-102,
    138-163) restated in oracle/crf_oracle.py against the outputs of the reference's own classes."""
    from oracle import crf_oracle as O
    z = _load("stereo_wrappers.npz")
    if "refiner_out" not in z.files:
        pytest.skip("stereo_wrappers.npz predates the Module forwards")
    t = lambda k: torch.from_numpy(z[k])
    g, s0 = torch.tensor(0.05), torch.tensor(0.0)
    out = O.crf_depth_refiner(t("refiner_logits"), t("refiner_img"), t("refiner_feats"),
                              t("refiner__projection__weight"), t("refiner__projection__bias"), g, s0, 2)
    np.testing.assert_allclose(out.numpy(), z["refiner_out"], rtol=0, atol=2e-5)
    un = {k[len("uncert__uncertainty_net__"):]: t(k) for k in z.files if k.startswith("uncert__uncertainty_net__")}
    conf = torch.exp(-O.uncertainty_net(t("refiner_img"), un))
    np.testing.assert_allclose(conf.numpy(), z["uncert_conf"], rtol=1e-5, atol=1e-6)
    out = O.crf_depth_refiner(t("refiner_logits"), t("refiner_img"), t("refiner_feats"),
                              t("uncert__projection__weight"), t("uncert__projection__bias"), g, s0, 2,
                              confidence=conf)
    np.testing.assert_allclose(out.numpy(), z["uncert_out"], rtol=0, atol=5e-5)
    out = O.crf_depth_upsampler(t("upsampler_disp"), t("upsampler_img"), g, s0, 2)
    np.testing.assert_allclose(out.numpy(), z["upsampler_out"], rtol=0, atol=2e-5)
    out = O.crf_depth_upsampler(t("upsampler_disp") * 1.7, t("upsampler_img"), g, s0, 2)
    np.testing.assert_allclose(out.numpy(), z["upsampler_out2"], rtol=0, atol=2e-5)


def test_depth_oracle_matches_reference_cost_volume_golden():
    """oracle/depth_oracle.py against crf.depth.disparity_badness itself (AD/9, SD/5, nprod/3)."""
    from oracle import depth_oracle as D
    z = _load("cost_volume_26x44.npz")
    left, right = z["left"], z["right"]
    np.testing.assert_allclose(D.disparity_badness(left, right), z["ad9"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(D.disparity_badness(left, right, 5, D.SD), z["sd5"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(D.disparity_badness(left, right, 3, D.nprod), z["np3"], rtol=0, atol=1e-12)
    assert np.array_equal(D.disparity_estimate(left, right), z["est"])


def test_fixture_readers_match_reference_golden(tmp_path):
    """crf/utils.py readers (host code of the package) on the byte streams the reference read."""
    import importlib
    U = importlib.import_module("depth_estimation_b200.crf.utils")
    z = _load("fixture_readers.npz")
    for name, reader, key in (("a.pgm", U.read_pgm, "pgm"), ("le.pfm", U.read_pfm, "pfm_le"),
                              ("be.pfm", U.read_pfm, "pfm_be"), ("c.png", U.read_image, "png")):
        path = tmp_path / name
        path.write_bytes(z[key + "_bytes"].tobytes())
        got = reader(str(path))
        assert got.shape == z[key].shape and got.dtype == z[key].dtype, name
        assert np.array_equal(got, z[key]), name
    # where the reference fails: 16-bit PGM (undefined `byteorder`), colour PFM, bad header
    wide = (np.arange(12) * 1000).astype(">u2").reshape(3, 4)
    (tmp_path / "w.pgm").write_bytes(b"P5\n4 3\n65535\n" + wide.tobytes())
    assert np.array_equal(U.read_pgm(str(tmp_path / "w.pgm")), wide)
    col = np.arange(24, dtype="<f4").reshape(2, 4, 3)
    (tmp_path / "c.pfm").write_bytes(b"PF\n4 2\n-1.0\n" + col.tobytes())
    assert np.array_equal(U.read_pfm(str(tmp_path / "c.pfm")), np.flip(col, 0).astype(np.float64))
    (tmp_path / "bad.pgm").write_bytes(b"P2\n1 1\n255\n0")
    with pytest.raises(ValueError):
        U.read_pgm(str(tmp_path / "bad.pgm"))

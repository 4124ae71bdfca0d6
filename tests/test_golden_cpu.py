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

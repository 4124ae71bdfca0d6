"""The C-ABI library loads on a CPU-only box and exports every symbol that
include/pl_lattice.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "pl_lattice.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pl_[a-z_0-9]+)\s*\(", text)))


def test_header_declares_expected_entry_points():
    names = _declared()
    for must in ("pl_filter", "pl_filter_host", "pl_lattice_create", "pl_lattice_apply",
                 "pl_lattice_destroy", "pl_meanfield", "pl_softmax_compat", "pl_lattice_export"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import depth_estimation_b200 as pkg
    lib = pkg._native.lib()
    for name in _declared():
        assert hasattr(lib, name), name
    assert set(pkg._native.EXPORTED_SYMBOLS) == set(_declared())
    assert lib.pl_version() == 1
    assert lib.pl_error_string(0) == b"ok"
    assert lib.pl_launch_count() == 0 or lib.pl_launch_count() > 0


def test_no_cpu_fallback_in_product_path():
    import torch
    import depth_estimation_b200 as pkg
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    src, ref = torch.zeros(4, 2), torch.zeros(4, 3)
    with pytest.raises(RuntimeError):
        pkg.lattice.filter(src, ref)
    with pytest.raises(RuntimeError):
        pkg.PermutohedralLattice(ref)


def test_product_never_imports_oracle():
    pkg_dir = os.path.join(ROOT, "depth-estimation_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("oracle/", "").lower() or f == "synthetic.py", \
                    os.path.join(dirpath, f)


def test_transposed_compat_cache_tracks_identity_and_version():
    """lattice._transposed(mu) (host logic, no device needed): one transposed copy per compatibility
    tensor, invalidated by an in-place change, never matched by another tensor of the same address."""
    import importlib
    import torch
    lat = importlib.import_module("depth_estimation_b200.lattice")
    lat._MU_T.clear()
    mu = torch.arange(12, dtype=torch.float32).reshape(3, 4)
    t1 = lat._transposed(mu)
    assert t1.shape == (4, 3) and t1.is_contiguous() and torch.equal(t1, mu.t())
    assert lat._transposed(mu) is t1                       # remembered
    mu.mul_(2.0)                                           # in-place change bumps the version counter
    t2 = lat._transposed(mu)
    assert t2 is not t1 and torch.equal(t2, mu.t())
    other = mu.clone()
    t3 = lat._transposed(other)
    assert t3 is not t2 and torch.equal(t3, other.t())
    # the cache keeps the tensors it has seen alive (so an id cannot be recycled) and stays small
    for _ in range(3 * lat._MU_T_MAX):
        lat._transposed(torch.rand(2, 2))
    assert len(lat._MU_T) <= lat._MU_T_MAX
    for key, (kept, _, _) in lat._MU_T.items():
        assert id(kept) == key
    # a tensor that requires grad is transposed detached
    g = torch.rand(2, 3, requires_grad=True)
    assert not lat._transposed(g).requires_grad

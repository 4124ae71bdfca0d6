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

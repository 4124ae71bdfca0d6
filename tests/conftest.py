import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle in the reference's own behaviour (hash-growth quirk included), which is
    what the product builds by default; identical to ``oracle_clean`` while M < 16383."""
    from oracle.lattice_oracle import Oracle
    return Oracle(faithful=True)


@pytest.fixture(scope="session")
def oracle_clean():
    """One vertex per distinct key (``PermutohedralLattice(ref, ref_grow_quirk=False)``)."""
    from oracle.lattice_oracle import Oracle
    return Oracle(faithful=False)


@pytest.fixture(scope="session")
def oracle_faithful():
    from oracle.lattice_oracle import Oracle
    return Oracle(faithful=True)


@pytest.fixture(scope="session")
def reference():
    from oracle.lattice_oracle import Reference
    if not Reference.available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return Reference()


@pytest.fixture(scope="session")
def native():
    """The CUDA library; GPU tests fail loudly if it is missing."""
    import torch
    assert torch.cuda.is_available(), "GPU test without a GPU"
    import depth_estimation_b200 as pkg
    pkg._native.lib()
    return pkg

"""B200-native permutohedral-lattice filter + dense-CRF mean-field loop.

Drop-in for the hot path of mfinzi/depth-estimation (crf/lattice/lite/lattice.cpp,
crf/gaussian_matrix.py lattice operators, crf/crf_module.py, the Module
forwards of crf/mb_stereo_crf.py).  Import as ``depth_estimation_b200`` (the
alias module at the repository root maps the name onto this directory).

    from depth_estimation_b200 import lattice            # lattice.filter(src, ref)
    from depth_estimation_b200.crf.gaussian_matrix import LatticeFilter, BatchedAdjacency
    from depth_estimation_b200.crf.crf_module import CRFasRNN, charb, ijrgbGuide
"""
from . import _native  # noqa: F401
from . import lattice  # noqa: F401
from .lattice import PermutohedralLattice, filter  # noqa: F401,A004

__all__ = ["lattice", "PermutohedralLattice", "filter"]

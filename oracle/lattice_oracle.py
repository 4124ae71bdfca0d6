"""ctypes bindings for the CPU oracle (TEST INFRASTRUCTURE, not product code).

Two back ends with the same Python surface:

* ``Oracle``     -- oracle/lattice_oracle.c, the plain-C restatement of
  crf/lattice/lite/permutohedral.h (built by ``make -C oracle``).
* ``Reference``  -- oracle/_ref/libref_lattice.so, the reference header itself
  compiled in place by oracle/build_ref.py (only where it was built).

All arrays are numpy, float32 / int16 / int8 / int32, C-contiguous.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(HERE, "liblattice_oracle.so")
_REF_SO = os.path.join(HERE, "_ref", "libref_lattice.so")

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i16p = np.ctypeslib.ndpointer(np.int16, flags="C_CONTIGUOUS")
_i8p = np.ctypeslib.ndpointer(np.int8, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def build_oracle(force=False):
    """Compile oracle/lattice_oracle.c if the .so is missing or stale."""
    src = os.path.join(HERE, "lattice_oracle.c")
    if force or not os.path.exists(_ORACLE_SO) or os.path.getmtime(_ORACLE_SO) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared",
                               "-o", _ORACLE_SO, src, "-lm"])
    return _ORACLE_SO


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class _Lattice:
    """Integer / float artefacts of one lattice build (numpy copies)."""

    def __init__(self, n, d, M, greedy, rank, bary, vid, keys, nbr, elevated=None):
        self.n, self.d, self.M = n, d, M
        self.greedy, self.rank, self.bary, self.vid = greedy, rank, bary, vid
        self.keys, self.nbr, self.elevated = keys, nbr, elevated


class Oracle:
    """The C restatement (oracle/lattice_oracle.c)."""

    kind = "port"

    def __init__(self, faithful=False):
        """faithful=True reproduces the reference's stale-hash-after-grow quirk
        (one possible orphan duplicate vertex per hash-table doubling, see
        lattice_oracle.c); faithful=False is the quirk-free lattice the product
        implements.  Identical whenever M < 16383 (no doubling)."""
        self.faithful = int(bool(faithful))
        lib = C.CDLL(build_oracle())
        lib.plo_build.restype = C.c_void_p
        lib.plo_build.argtypes = [_f32p, C.c_int, C.c_int, C.c_int]
        lib.plo_num_grows.argtypes = [C.c_void_p]
        lib.plo_free.argtypes = [C.c_void_p]
        lib.plo_apply.argtypes = [C.c_void_p, _f32p, C.c_int, _f32p]
        lib.plo_filter.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        lib.plo_num_vertices.argtypes = [C.c_void_p]
        lib.plo_scale_factors.argtypes = [C.c_int, _f32p]
        for name, ty in [("plo_elevated", C.c_float), ("plo_greedy", C.c_int16), ("plo_rank", C.c_int8),
                         ("plo_bary", C.c_float), ("plo_vid", C.c_int32), ("plo_keys", C.c_int16),
                         ("plo_nbr", C.c_int32)]:
            fn = getattr(lib, name)
            fn.restype = C.POINTER(ty)
            fn.argtypes = [C.c_void_p]
        self.lib = lib

    def scale_factors(self, d):
        out = np.empty(d, np.float32)
        self.lib.plo_scale_factors(d, out)
        return out

    def filter(self, src, ref):
        """lattice.cpp:6-10: filter(src[n,C], ref[n,d]) -> [n,C]."""
        src, ref = _f32(src), _f32(ref)
        n, c = src.shape
        assert ref.shape[0] == n
        out = np.empty((n, c), np.float32)
        self.lib.plo_filter(src, ref, n, c, ref.shape[1], self.faithful, out)
        return out

    def build(self, ref):
        ref = _f32(ref)
        n, d = ref.shape
        h = self.lib.plo_build(ref, n, d, self.faithful)
        return _Handle(self.lib, h, n, d)

    def state(self, ref):
        h = self.build(ref)
        try:
            return h.state()
        finally:
            h.close()


class _Handle:
    def __init__(self, lib, h, n, d):
        self.lib, self.h, self.n, self.d = lib, h, n, d
        self.M = lib.plo_num_vertices(h)
        self.ngrow = lib.plo_num_grows(h)

    def _arr(self, fn, shape, dtype):
        ptr = getattr(self.lib, fn)(self.h)
        count = int(np.prod(shape))
        if count == 0:
            return np.zeros(shape, dtype)
        return np.ctypeslib.as_array(ptr, shape=(count,)).astype(dtype).reshape(shape).copy()

    def state(self):
        n, d, M = self.n, self.d, self.M
        return _Lattice(
            n, d, M,
            greedy=self._arr("plo_greedy", (n, d + 1), np.int16),
            rank=self._arr("plo_rank", (n, d + 1), np.int8),
            bary=self._arr("plo_bary", (n, d + 1), np.float32),
            vid=self._arr("plo_vid", (n, d + 1), np.int32),
            keys=self._arr("plo_keys", (M, d), np.int16),
            nbr=self._arr("plo_nbr", (d + 1, M, 2), np.int32),
            elevated=self._arr("plo_elevated", (n, d + 1), np.float32),
        )

    def apply(self, src):
        src = _f32(src)
        assert src.shape[0] == self.n
        out = np.empty_like(src)
        self.lib.plo_apply(self.h, src, src.shape[1], out)
        return out

    def close(self):
        if self.h:
            self.lib.plo_free(self.h)
            self.h = None

    def __del__(self):
        self.close()


class Reference:
    """The reference header itself (oracle/_ref/libref_lattice.so)."""

    kind = "reference"

    @staticmethod
    def available():
        return os.path.exists(_REF_SO)

    def __init__(self):
        if not os.path.exists(_REF_SO):
            raise FileNotFoundError(f"{_REF_SO} not built (run python oracle/build_ref.py where /root/reference exists)")
        import torch  # noqa: F401  (libref_lattice.so links libtorch; load it first)
        lib = C.CDLL(_REF_SO)
        lib.ref_filter.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, _f32p]
        lib.ref_state.restype = C.c_int
        lib.ref_state.argtypes = [_f32p, C.c_int, C.c_int, _i16p, _i8p, _f32p, _i32p, _i16p, C.c_int]
        lib.ref_neighbours.restype = C.c_int
        lib.ref_neighbours.argtypes = [_f32p, C.c_int, C.c_int, _i32p, C.c_int]
        self.lib = lib

    def filter(self, src, ref):
        src, ref = _f32(src), _f32(ref)
        n, c = src.shape
        out = np.empty((n, c), np.float32)
        self.lib.ref_filter(src, ref, n, c, ref.shape[1], out)
        return out

    def state(self, ref):
        ref = _f32(ref)
        n, d = ref.shape
        cap = n * (d + 1)
        greedy = np.empty((n, d + 1), np.int16)
        rank = np.empty((n, d + 1), np.int8)
        bary = np.empty((n, d + 1), np.float32)
        vid = np.empty((n, d + 1), np.int32)
        keys = np.empty((cap, d), np.int16)
        M = self.lib.ref_state(ref, n, d, greedy, rank, bary, vid, keys, cap)
        nbr = np.empty((d + 1, M, 2), np.int32)
        M2 = self.lib.ref_neighbours(ref, n, d, nbr, M)
        assert M2 == M
        return _Lattice(n, d, M, greedy, rank, bary, vid, keys[:M].copy(), nbr)

"""ctypes binding of libpl_lattice.so (the C ABI in include/pl_lattice.h).

PyTorch supplies device memory and streams; every computation goes through the
C ABI below.  There is no fallback: if the library is missing or no sm_100 GPU
is present the calls raise.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpl_lattice.so")

PL_OK = 0
PL_BUILD_DEFAULT = 0
PL_BUILD_REF_GROW_QUIRK = 1
PL_APPLY_DEFAULT = 0
PL_APPLY_EXACT_SLICE = 1
PL_APPLY_FLAT_KERNELS = 2
PL_MEANFIELD_UNFUSED = 4
PL_MEANFIELD_NO_TENSOR = 8
PL_MEANFIELD_TILE_KERNEL = 16
(PL_EXPORT_GREEDY, PL_EXPORT_RANK, PL_EXPORT_BARY, PL_EXPORT_VERTEX, PL_EXPORT_KEYS,
 PL_EXPORT_NEIGHBOURS) = range(6)

_lib = None

_SIGNATURES = {
    # name: (restype, argtypes)
    "pl_lattice_create": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_uint32, C.c_void_p,
                                    C.POINTER(C.c_void_p)]),
    "pl_lattice_create_ijrgb": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float,
                                          C.c_uint32, C.c_void_p, C.POINTER(C.c_void_p)]),
    "pl_lattice_destroy": (None, [C.c_void_p]),
    "pl_lattice_num_points": (C.c_int64, [C.c_void_p]),
    "pl_lattice_dim": (C.c_int, [C.c_void_p]),
    "pl_lattice_num_vertices": (C.c_int64, [C.c_void_p]),
    "pl_lattice_device_bytes": (C.c_int64, [C.c_void_p]),
    "pl_lattice_apply": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int,
                                   C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_uint32,
                                   C.c_void_p]),
    "pl_filter": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int,
                            C.c_void_p]),
    "pl_filter_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int,
                                 C.c_void_p]),
    "pl_softmax_compat": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                    C.c_int, C.c_float, C.c_void_p]),
    "pl_softmax_compat_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                            C.c_int, C.c_float, C.c_void_p]),
    "pl_softmax_compat_ex": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                       C.c_int, C.c_float, C.c_uint32, C.c_void_p]),
    "pl_softmax_compat_blocked": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64,
                                            C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p]),
    "pl_softmax_compat_peers": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_int64, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p]),
    "pl_softmax_compat_shards": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int, C.c_void_p,
                                           C.c_int64, C.c_int64, C.c_int, C.c_float, C.c_void_p]),
    "pl_peer_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p), C.c_void_p]),
    "pl_peer_open": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "pl_peer_close": (C.c_int, [C.c_void_p]),
    "pl_peer_free": (C.c_int, [C.c_void_p]),
    "pl_meanfield": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                               C.c_int, C.c_float, C.c_uint32, C.c_void_p]),
    "pl_meanfield_depth": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_uint32,
                                     C.c_void_p]),
    "pl_softargmax": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_float, C.c_void_p,
                                C.c_void_p, C.c_void_p]),
    "pl_cost_volume": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                 C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "pl_lattice_apply_dot": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64,
                                       C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_float, C.c_void_p]),
    "pl_lattice_apply_dot_multi": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64,
                                             C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_float,
                                             C.c_void_p]),
    "pl_meanfield_backward_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_int64, C.c_int, C.c_float, C.c_void_p]),
    "pl_softmax_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int,
                                      C.c_float, C.c_void_p]),
    "pl_nchw_to_nlc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_float,
                                 C.c_void_p]),
    "pl_nlc_to_nchw": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_float,
                                 C.c_void_p]),
    "pl_lattice_export": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "pl_profile": (None, [C.c_int]),
    "pl_profile_report": (C.c_int, [C.c_char_p, C.c_int]),
    "pl_launch_count": (C.c_uint64, []),
    "pl_debug_tc_timeouts": (C.c_int, []),
    "pl_version": (C.c_int, []),
    "pl_error_string": (C.c_char_p, [C.c_int]),
    "pl_last_error": (C.c_char_p, []),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def lib():
    """Loads libpl_lattice.so once.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python depth-estimation_b200/csrc/build.py` "
                "(there is no CPU or PyTorch fallback for the lattice path)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


class LatticeError(RuntimeError):
    pass


def check(code):
    if code != PL_OK:
        l = lib()
        raise LatticeError(f"libpl_lattice: {l.pl_error_string(code).decode()}: "
                           f"{l.pl_last_error().decode()}")


def stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda_f32(t, name, ndim=None):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise LatticeError(f"{name} must be a CUDA tensor (the lattice path has no CPU fallback)")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32, got {t.dtype}")
    if ndim is not None and t.dim() != ndim:
        raise ValueError(f"{name} must be {ndim}-D, got shape {tuple(t.shape)}")


def profile(enable):
    lib().pl_profile(1 if enable else 0)


def profile_report():
    """{kernel name: (launches, total_ms)} since profiling was switched on; clears."""
    buf = C.create_string_buffer(1 << 16)
    lib().pl_profile_report(buf, len(buf))
    out = {}
    for line in buf.value.decode().splitlines():
        name, cnt, ms = line.split("\t")
        out[name.strip("()")] = (int(cnt), float(ms))
    return out


def launch_count():
    return int(lib().pl_launch_count())

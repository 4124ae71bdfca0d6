"""Drop-in for the reference's native module ``lattice`` (crf/lattice/lite/lattice.cpp:6-16).

The reference JIT-builds a pybind module exporting one function,
``filter(src[n,C], ref[n,d]) -> [n,C]`` (crf/gaussian_matrix.py:15-16 binds it
as ``latticefilter``).  This module exports the same ``filter`` on top of the
C ABI (include/pl_lattice.h) plus the handle the reference lacks:
``PermutohedralLattice(ref)`` builds the lattice once, ``.filter(src)`` applies
it any number of times (every mean-field iteration and the backward pass reuse
one lattice; the reference rebuilds it per call, permutohedral.h:208,236-238).

Tensors: fp32.  CUDA tensors stay on the device.  CPU tensors (what the
reference op takes) are copied to the current CUDA device, filtered there and
copied back -- there is no CPU implementation in this package.
"""
import ctypes as C

import torch

from . import _native as N


_MU_T = {}       # id(mu) -> (mu, version, transposed copy); holds mu itself, so its id cannot be reused
_MU_T_MAX = 4


def _transposed(mu):
    """``mu.t().contiguous()`` for the C ABI (which takes Mu^T), remembered per compatibility tensor:
    a streaming loop passes the same Mu every frame, and the transposition is a kernel launch plus
    host time that sits on the critical path between the lattice build and the first iteration.
    The entry keeps a reference to ``mu`` (the key is its identity, the in-place version counter
    invalidates it), so a recycled address or id can never match."""
    ent = _MU_T.get(id(mu))
    if ent is not None and ent[0] is mu and ent[1] == mu._version:
        return ent[2]
    mu_t = mu.detach().t().contiguous()
    if len(_MU_T) >= _MU_T_MAX:
        _MU_T.pop(next(iter(_MU_T)))
    _MU_T[id(mu)] = (mu, mu._version, mu_t)
    return mu_t


class PermutohedralLattice:
    """Cached lattice for one feature matrix ``ref[n, d]`` (CUDA, fp32).

    Mirrors the reference class of the same name (permutohedral.h:186-578):
    construction = its constructor + the position-dependent half of splat()
    for all points; ``filter`` = splat values -> blur -> slice.
    """

    def __init__(self, ref, ref_grow_quirk=True):
        """``ref_grow_quirk`` (default on): reproduce the reference's hash-table growth behaviour
        (permutohedral.h:59-62,101-155: the lookup that doubles the table probes from a stale
        slot and can add a duplicate vertex), so vertex count, ids, keys, neighbours and filtered
        values equal the reference's for every M; off = one vertex per distinct key.  The two are
        the same lattice while M < 16383."""
        N.require_cuda_f32(ref, "ref", 2)
        ref = ref.contiguous()
        self.device = ref.device
        self.n, self.d = int(ref.shape[0]), int(ref.shape[1])
        self._h = C.c_void_p()
        flags = N.PL_BUILD_REF_GROW_QUIRK if ref_grow_quirk else N.PL_BUILD_DEFAULT
        with torch.cuda.device(self.device):
            N.check(N.lib().pl_lattice_create(C.c_void_p(ref.data_ptr()), self.n, self.d, flags,
                                              N.stream_ptr(self.device), C.byref(self._h)))
        self.num_vertices = int(N.lib().pl_lattice_num_vertices(self._h))

    @classmethod
    def from_rgb(cls, rgb, s_ij=0.1, s_rgb=0.1, ref_grow_quirk=True):
        """Lattice of ``ijrgbGuide(s_ij, s_rgb)(rgb[None])`` (crf_module.py:106-114) for one image
        ``rgb[3, h, w]``: the features are evaluated inside the elevation kernel, bit for bit
        what the module computes, and never written to memory."""
        N.require_cuda_f32(rgb, "rgb", 3)
        if rgb.shape[0] != 3:
            raise ValueError(f"rgb must be [3, h, w], got {tuple(rgb.shape)}")
        rgb = rgb.contiguous()
        self = cls.__new__(cls)
        self.device = rgb.device
        h, w = int(rgb.shape[1]), int(rgb.shape[2])
        self.n, self.d = h * w, 5
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            N.check(N.lib().pl_lattice_create_ijrgb(
                C.c_void_p(rgb.data_ptr()), h, w, float(s_ij), float(s_rgb),
                N.PL_BUILD_REF_GROW_QUIRK if ref_grow_quirk else N.PL_BUILD_DEFAULT,
                N.stream_ptr(self.device), C.byref(self._h)))
        self.num_vertices = int(N.lib().pl_lattice_num_vertices(self._h))
        return self

    # -- value path ---------------------------------------------------------
    def filter(self, src, sub=None, add=None, exact_slice=False, out=None, flat=False):
        """``filter(src) - sub + add`` for ``src[n, C]`` (row stride may exceed C).

        ``sub=src`` is the adjacency ``W(x) = K x - x`` (gaussian_matrix.py:302,352);
        ``add=E0`` fuses the mean-field update (crf_module.py:101).
        """
        self._check_rows(src, "src")
        src = _rowmajor(src)
        c = int(src.shape[1])
        if out is None:
            out = torch.empty((self.n, c), dtype=torch.float32, device=self.device)
        else:
            self._check_rows(out, "out")
            if out.shape[1] != c or out.stride(1) != 1:
                raise ValueError("out must be [n, C] with unit channel stride")
        ptrs, strides = [], []
        keep = []   # row-major temporaries must outlive the launch (their storage is the operand)
        for name, t in (("sub", sub), ("add", add)):
            if t is None:
                ptrs.append(C.c_void_p(0))
                strides.append(0)
            else:
                self._check_rows(t, name)
                if t.shape[1] != c:
                    raise ValueError(f"{name} must have {c} channels")
                t = _rowmajor(t)
                keep.append(t)
                ptrs.append(C.c_void_p(t.data_ptr()))
                strides.append(t.stride(0))
        flags = (N.PL_APPLY_EXACT_SLICE if exact_slice else 0) | (N.PL_APPLY_FLAT_KERNELS if flat else 0)
        with torch.cuda.device(self.device):
            N.check(N.lib().pl_lattice_apply(
                self._h, C.c_void_p(src.data_ptr()), src.stride(0), C.c_void_p(out.data_ptr()),
                out.stride(0), c, ptrs[0], strides[0], ptrs[1], strides[1], flags,
                N.stream_ptr(self.device)))
        del keep
        return out

    __call__ = filter

    def filter_dot_multi(self, src, other, out, rowscale, coef=1.0):
        """``out[p, j] += coef * <filter(src * rowscale[:, j, None])[p], other[p]>`` for all the
        J = 5 columns of ``rowscale [n, 5]`` in one pass (d = 5 lattices; gaussian_matrix.py:450-463)."""
        self._check_rows(src, "src")
        self._check_rows(other, "other")
        src, other = _rowmajor(src), _rowmajor(other)
        c = int(src.shape[1])
        J = int(rowscale.shape[1])
        if (other.shape[1] != c or tuple(out.shape) != (self.n, J) or tuple(rowscale.shape) != (self.n, J)
                or rowscale.stride(1) != 1 or out.stride(1) != 1 or rowscale.dtype != torch.float32):
            raise ValueError("filter_dot_multi: other [n, C], rowscale / out [n, J] with unit column stride")
        with torch.cuda.device(self.device):
            N.check(N.lib().pl_lattice_apply_dot_multi(
                self._h, C.c_void_p(src.data_ptr()), src.stride(0), c, C.c_void_p(rowscale.data_ptr()),
                rowscale.stride(0), J, C.c_void_p(other.data_ptr()), other.stride(0),
                C.c_void_p(out.data_ptr()), out.stride(0), float(coef), N.stream_ptr(self.device)))
        return out

    def filter_dot(self, src, other, out, rowscale=None, coef=1.0):
        """``out[p] += coef * <filter(src * rowscale[:, None])[p], other[p]>`` without materialising
        the scaled source or the filtered rows (the guide gradient, gaussian_matrix.py:450-463).
        ``out`` and ``rowscale`` are 1-D views with any stride (columns of [n, d] matrices)."""
        self._check_rows(src, "src")
        self._check_rows(other, "other")
        src, other = _rowmajor(src), _rowmajor(other)
        c = int(src.shape[1])
        if other.shape[1] != c or out.shape[0] != self.n or out.dim() != 1:
            raise ValueError("filter_dot: other must be [n, C] and out [n]")
        rs_ptr, rs_stride = C.c_void_p(0), 0
        if rowscale is not None:
            if rowscale.dim() != 1 or rowscale.shape[0] != self.n or rowscale.dtype != torch.float32:
                raise ValueError("filter_dot: rowscale must be float32 [n]")
            rs_ptr, rs_stride = C.c_void_p(rowscale.data_ptr()), rowscale.stride(0)
        with torch.cuda.device(self.device):
            N.check(N.lib().pl_lattice_apply_dot(
                self._h, C.c_void_p(src.data_ptr()), src.stride(0), c, rs_ptr, rs_stride,
                C.c_void_p(other.data_ptr()), other.stride(0), C.c_void_p(out.data_ptr()), out.stride(0),
                float(coef), N.stream_ptr(self.device)))
        return out


    def mean_field(self, e0, mu, niters, scale=1.0, exact_slice=False, flat=False, unfused=False,
                   no_tensor=False, tile_kernel=False):
        """Pixel-major CRFasRNN inference loop (crf_module.py:97-103).

        ``e0[n, L]`` unary energies, ``mu[L, L]`` compatibility matrix
        (``Mu[l][k]``, crf_module.py:74); returns E of the last iteration.
        """
        self._check_rows(e0, "e0")
        e0 = e0.contiguous()
        L = int(e0.shape[1])
        N.require_cuda_f32(mu, "mu", 2)
        if tuple(mu.shape) != (L, L):
            raise ValueError(f"mu must be [{L}, {L}]")
        mu_t = _transposed(mu)
        out = torch.empty_like(e0)
        work = torch.empty_like(e0)
        flags = ((N.PL_APPLY_EXACT_SLICE if exact_slice else 0) | (N.PL_APPLY_FLAT_KERNELS if flat else 0)
                 | (N.PL_MEANFIELD_UNFUSED if unfused else 0)
                 | (N.PL_MEANFIELD_NO_TENSOR if no_tensor else 0)
                 | (N.PL_MEANFIELD_TILE_KERNEL if tile_kernel else 0))
        with torch.cuda.device(self.device):
            N.check(N.lib().pl_meanfield(
                self._h, C.c_void_p(e0.data_ptr()), C.c_void_p(mu_t.data_ptr()),
                C.c_void_p(out.data_ptr()), C.c_void_p(work.data_ptr()), L, int(niters),
                float(scale), flags, N.stream_ptr(self.device)))
        return out

    def mean_field_depth(self, e0, mu, niters, scale=1.0, labels=None, want_depth=True,
                         want_label=False, unfused=False, no_tensor=False, tile_kernel=False):
        """``mean_field`` followed by the read-out, fused into the last slice: soft-argmax depth
        ``sum_l softmax(-E)_l * labels[l]`` (mb_stereo_crf.py:62-66) and / or the arg-min label.
        Returns ``(depth[n] or None, label[n] int32 or None)``; E itself is not materialised."""
        self._check_rows(e0, "e0")
        e0 = e0.contiguous()
        L = int(e0.shape[1])
        N.require_cuda_f32(mu, "mu", 2)
        if tuple(mu.shape) != (L, L):
            raise ValueError(f"mu must be [{L}, {L}]")
        mu_t = _transposed(mu)
        if labels is not None:
            N.require_cuda_f32(labels, "labels")
            labels = labels.reshape(-1).contiguous()
            if labels.numel() != L:
                raise ValueError(f"labels must have {L} entries")
        depth = torch.empty((self.n,), dtype=torch.float32, device=self.device) if want_depth else None
        label = torch.empty((self.n,), dtype=torch.int32, device=self.device) if want_label else None
        if depth is None and label is None:
            raise ValueError("nothing to read out")
        work = torch.empty_like(e0)
        flags = ((N.PL_MEANFIELD_UNFUSED if unfused else 0)
                 | (N.PL_MEANFIELD_NO_TENSOR if no_tensor else 0)
                 | (N.PL_MEANFIELD_TILE_KERNEL if tile_kernel else 0))
        ptr = lambda t: C.c_void_p(t.data_ptr() if t is not None else 0)
        with torch.cuda.device(self.device):
            N.check(N.lib().pl_meanfield_depth(
                self._h, ptr(e0), ptr(mu_t), ptr(depth), ptr(label), ptr(work), ptr(labels), L,
                int(niters), float(scale), flags, N.stream_ptr(self.device)))
        return depth, label

    # -- inspection (parity tests) -------------------------------------------
    def export(self):
        """Integer / float artefacts as torch tensors (see PL_EXPORT_* in pl_lattice.h)."""
        n, d, M = self.n, self.d, self.num_vertices
        dev = self.device
        out = {
            "greedy": (N.PL_EXPORT_GREEDY, torch.empty((n, d + 1), dtype=torch.int16, device=dev)),
            "rank": (N.PL_EXPORT_RANK, torch.empty((n, d + 1), dtype=torch.int8, device=dev)),
            "bary": (N.PL_EXPORT_BARY, torch.empty((n, d + 1), dtype=torch.float32, device=dev)),
            "vid": (N.PL_EXPORT_VERTEX, torch.empty((n, d + 1), dtype=torch.int32, device=dev)),
            "keys": (N.PL_EXPORT_KEYS, torch.empty((M, d), dtype=torch.int16, device=dev)),
            "nbr": (N.PL_EXPORT_NEIGHBOURS,
                    torch.empty((d + 1, M, 2), dtype=torch.int32, device=dev)),
        }
        with torch.cuda.device(dev):
            for what, t in out.values():
                N.check(N.lib().pl_lattice_export(self._h, what, C.c_void_p(t.data_ptr()),
                                                  N.stream_ptr(dev)))
        return {k: t for k, (_, t) in out.items()}

    @property
    def device_bytes(self):
        return int(N.lib().pl_lattice_device_bytes(self._h))

    def _check_rows(self, t, name):
        N.require_cuda_f32(t, name, 2)
        if t.device != self.device:
            raise ValueError(f"{name} is on {t.device}, lattice on {self.device}")
        if t.shape[0] != self.n:
            # same message as the reference's assert (gaussian_matrix.py:429-430)
            raise AssertionError("Incompatible shapes {}, and {}".format(tuple(t.shape),
                                                                        (self.n, self.d)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            N.lib().pl_lattice_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _rowmajor(t):
    """[n, C] view with unit channel stride and a 16-byte friendly row stride."""
    if t.stride(1) != 1 or t.stride(0) < t.shape[1]:
        t = t.contiguous()
    return t


def filter(src, ref):  # noqa: A001 - the reference's exported name
    """``lattice.filter(src[n,C], ref[n,d]) -> [n,C]`` (lattice.cpp:6-10).

    Output lives where ``src`` lives.  CPU inputs are staged through the GPU.
    """
    if src.dim() != 2 or ref.dim() != 2:
        raise ValueError("filter expects src[n,C] and ref[n,d]")
    assert src.shape[0] == ref.shape[0], \
        "Incompatible shapes {}, and {}".format(tuple(src.shape), tuple(ref.shape))
    if src.dtype != torch.float32 or ref.dtype != torch.float32:
        raise TypeError("filter expects float32 tensors (permutohedral.h:19 float_type)")
    n, c, d = int(src.shape[0]), int(src.shape[1]), int(ref.shape[1])
    if not src.is_cuda:
        # the reference's own call shape: host buffers in, host buffer out
        if not torch.cuda.is_available():
            raise N.LatticeError("lattice.filter needs a CUDA device (no CPU fallback)")
        src_c, ref_c = src.contiguous(), ref.cpu().contiguous()
        out = torch.empty_like(src_c, pin_memory=src_c.is_pinned())
        N.check(N.lib().pl_filter_host(C.c_void_p(src_c.data_ptr()), C.c_void_p(ref_c.data_ptr()),
                                       C.c_void_p(out.data_ptr()), n, c, d,
                                       N.stream_ptr(torch.cuda.current_device())))
        return out
    N.require_cuda_f32(ref, "ref", 2)
    src_c, ref_c = src.contiguous(), ref.contiguous()
    out = torch.empty_like(src_c)
    with torch.cuda.device(src.device):
        N.check(N.lib().pl_filter(C.c_void_p(src_c.data_ptr()), C.c_void_p(ref_c.data_ptr()),
                                  C.c_void_p(out.data_ptr()), n, c, d, N.stream_ptr(src.device)))
    return out

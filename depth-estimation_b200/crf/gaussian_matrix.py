"""Lattice pairwise operators -- mirror of the lattice half of crf/gaussian_matrix.py.

Same names, argument meaning and error behaviour as the reference:
``latticefilter`` (:16), ``LatticeFilter`` (:423-468), ``BatchedLatticeFilter``
(:379-421), ``batched_filter`` (:370-377), ``LatticeGaussian`` (:292-303),
``RbfLaplacian`` / ``RbfLaplacianC`` (:305-338), ``BatchedAdjacency`` (:341-352).

Differences that are the point of this package:
* the filter runs on the GPU through libpl_lattice (no CPU path);
* the lattice of a feature matrix is built once and reused by every call that
  passes the same tensor (forward, backward, every mean-field iteration) --
  the reference rebuilds it inside every ``latticefilter`` call;
* the batch is handled in one process on one stream instead of ``mp.Pool``
  (``num_threads`` is accepted and ignored);
* backward does not materialise the reference's ``[g, g*ref, src, src*ref]``
  concatenation (n x 2L(1+d)) in one piece when it is large: it is filtered in
  column blocks against the same cached lattice.  Same arithmetic per column.
"""
import torch
import torch.nn as nn
from torch.autograd import Function

from .. import lattice as _lattice
from ..lattice import PermutohedralLattice

latticefilter = _lattice.filter  # gaussian_matrix.py:16

# ---- lattice cache ----------------------------------------------------------
# One lattice per live feature tensor.  An entry PINS the tensor it was built
# from, so its storage cannot be freed and handed to a different tensor while
# the entry lives; the key also carries the version counter, so an in-place
# update of the features (through any view) invalidates it.  Only DETACHED
# values are stored (flattened guides without autograd history, lattices); the
# differentiable flattening lives outside the cache (_FlattenGuide below), so a
# no-grad call can never hand a grad-less guide to a later training call.
#
# The cache is a small LRU: a frame holds two entries (flattened guide and its
# lattice) per image, and a 2.9 MP lattice pins ~0.3 GB, so the default keeps
# only a few frames alive.  set_lattice_cache_limit() raises it for callers
# that revisit many guides (e.g. several Laplacians side by side).
_CACHE = {}
_CACHE_MAX = 8
_CACHE_LOCK = __import__("threading").RLock()   # callers may drive several streams from several threads


def _key(t):
    return (t.data_ptr(), tuple(t.shape), tuple(t.stride()), t._version, str(t.device))


def _cached(t, build):
    key = _key(t)
    with _CACHE_LOCK:
        hit = _CACHE.pop(key, None)
        if hit is not None:
            _CACHE[key] = hit                      # most recently used goes last
            return hit[1]
        while len(_CACHE) >= _CACHE_MAX:
            old = _CACHE.pop(next(iter(_CACHE)))[1]
            if isinstance(old, PermutohedralLattice):
                old.close()
    value = build()                                # outside the lock: a build waits for the device
    with _CACHE_LOCK:
        _CACHE[key] = (t, value)
    return value


def lattice_for(ref):
    """Cached ``PermutohedralLattice`` of ``ref[n, d]`` (CUDA fp32)."""
    return _cached(ref, lambda: PermutohedralLattice(ref.detach()))


def flat_guide(guide_imgs):
    """Detached pixel-major copy ``[B, n, d]`` of a guide batch ``[B, d, h, w]``
    (gaussian_matrix.py:348-349), cached per guide tensor so its lattices are found again on
    every mean-field iteration and in the backward pass."""
    bs, d = guide_imgs.shape[:2]
    return _cached(guide_imgs,
                   lambda: guide_imgs.detach().reshape(bs, d, -1).permute(0, 2, 1).contiguous())


class _FlattenGuide(Function):
    """Differentiable ``guide.reshape(B, d, -1).permute(0, 2, 1)`` whose forward value is a view of
    the cached detached copy (same storage, shape and version counter -> the same cache key, so the
    lattice built for it is reused by every iteration), and whose backward is the inverse permute."""

    @staticmethod
    def forward(ctx, guide_imgs):
        ctx.shape = tuple(guide_imgs.shape)
        flat = flat_guide(guide_imgs)
        return flat.view_as(flat)

    @staticmethod
    def backward(ctx, grad):
        return grad.permute(0, 2, 1).reshape(ctx.shape)


def set_lattice_cache_limit(entries):
    """Bound the cache (least recently used entries are destroyed first).  A frame holds two
    entries per image (the flattened guide and its lattice); a streaming caller that keeps k
    frames in flight wants 2k so device memory is returned to the pool as it goes."""
    global _CACHE_MAX
    _CACHE_MAX = max(2, int(entries))
    while len(_CACHE) > _CACHE_MAX:
        old = _CACHE.pop(next(iter(_CACHE)))[1]
        if isinstance(old, PermutohedralLattice):
            old.close()


def forget(guide_imgs):
    """Drop (and free) what the cache holds for one guide batch ``[B, d, h, w]``: its flattened
    copy and the lattices built from it.  For streaming callers: call once the frame's work
    has been enqueued (the lattice's memory is released in stream order)."""
    with _CACHE_LOCK:
        hit = _CACHE.pop(_key(guide_imgs), None)
    if hit is None:
        return
    flat_refs = hit[1]
    if isinstance(flat_refs, torch.Tensor) and flat_refs.dim() == 3:
        for b in range(flat_refs.shape[0]):
            with _CACHE_LOCK:
                lat = _CACHE.pop(_key(flat_refs[b]), None)
            if lat is not None and isinstance(lat[1], PermutohedralLattice):
                lat[1].close()


def clear_lattice_cache():
    while _CACHE:
        old = _CACHE.popitem()[1][1]
        if isinstance(old, PermutohedralLattice):
            old.close()


def _filter_cached(src, ref):
    return lattice_for(ref).filter(src.detach())


def _grad_reference(lat, src, ref, g):
    """gaussian_matrix.py:450-463:
        grad_ref = -2 * sum_L ( sf*wg - src*wgf + gf*ws - g*wsf ),
    sf = src (x) ref, gf = g (x) ref, w* = filter of the same.  The four filtered
    blocks are produced by one wide filter call in the reference; here the
    L*d-wide blocks are filtered in slabs of labels to bound memory."""
    n, L = src.shape
    d = ref.shape[1]
    wg = lat.filter(g)
    if (L % 4 == 0 and src.is_cuda and src.dtype == torch.float32 and g.dtype == torch.float32
            and ref.dtype == torch.float32 and (lat.num_vertices + 1) * L * 4 <= (2 << 30)):
        # the filtered outer products are only used inside inner products with the rows of src / g:
        # per feature j, splat (g * ref_j) resp. (src * ref_j) with the factor folded into the entry
        # weights and contract the slice with the other matrix on the spot (pl_lattice_apply_dot);
        # neither the n x 2Ld concatenation nor its filtered copy exists
        srcc, gc = src.contiguous(), g.contiguous()
        t = (srcc * wg).sum(1)                       # sum_L src * W g
        lat.filter_dot(srcc, gc, t)                  # + sum_L g * W src
        acc = (t[:, None] * ref).contiguous()        # [n, d]
        if d == 5 and ref.stride(1) == 1 and lat.num_vertices < (1 << 24):
            # all five features in one pass per side: source rows staged once for the five scaled
            # splats, rows of the other matrix read once for the five contractions
            lat.filter_dot_multi(gc, srcc, acc, ref, coef=-1.0)
            lat.filter_dot_multi(srcc, gc, acc, ref, coef=-1.0)
            return -2 * acc, wg
        for j in range(d):
            lat.filter_dot(gc, srcc, acc[:, j], rowscale=ref[:, j], coef=-1.0)
            lat.filter_dot(srcc, gc, acc[:, j], rowscale=ref[:, j], coef=-1.0)
        return -2 * acc, wg
    ws = lat.filter(src)
    # terms without the outer products: sum_L (src*wg + g*ws) * ref
    acc = (src * wg + g * ws).sum(1, keepdim=True) * ref  # [n, d]
    slab = max(1, min(L, (1 << 28) // max(1, n * d * 4)))
    for l0 in range(0, L, slab):
        l1 = min(L, l0 + slab)
        gs, ss = g[:, l0:l1], src[:, l0:l1]
        both = torch.cat([(gs[:, :, None] * ref[:, None, :]).reshape(n, -1),
                          (ss[:, :, None] * ref[:, None, :]).reshape(n, -1)], dim=1)
        w = lat.filter(both)
        k = (l1 - l0) * d
        wgf = w[:, :k].view(n, l1 - l0, d)
        wsf = w[:, k:].view(n, l1 - l0, d)
        acc = acc - (ss[:, :, None] * wgf + gs[:, :, None] * wsf).sum(1)
    return -2 * acc, wg


class LatticeFilter(Function):
    """gaussian_matrix.py:423-468."""

    @staticmethod
    def forward(ctx, source, reference):
        assert source.shape[0] == reference.shape[0], \
            "Incompatible shapes {}, and {}".format(source.shape, reference.shape)
        ctx.save_for_backward(source, reference)
        return _filter_cached(source, reference)

    @staticmethod
    def backward(ctx, grad_output):
        with torch.no_grad():
            src, ref = ctx.saved_tensors
            g = grad_output.contiguous()
            grad_source = grad_reference = None
            lat = lattice_for(ref)
            if ctx.needs_input_grad[0] and not ctx.needs_input_grad[1]:
                grad_source = lat.filter(g)  # "Matrix is symmetric", :446
            if ctx.needs_input_grad[1]:
                grad_reference, wg = _grad_reference(lat, src.contiguous(), ref, g)
                if ctx.needs_input_grad[0]:
                    grad_source = wg
        return grad_source, grad_reference


def batched_filter(flat_srcs, flat_refs, num_threads=None):
    """gaussian_matrix.py:370-377 without the process pool: one lattice per batch item."""
    return torch.stack([_filter_cached(s, r) for s, r in zip(flat_srcs, flat_refs)])


class BatchedLatticeFilter(Function):
    """gaussian_matrix.py:379-421."""

    @staticmethod
    def forward(ctx, flat_srcs, flat_refs, num_threads):
        assert flat_srcs.shape[:2] == flat_refs.shape[:2], \
            "Incompatible shapes {}, and {}".format(flat_srcs.shape, flat_refs.shape)
        ctx.save_for_backward(flat_srcs, flat_refs)
        ctx.num_threads = num_threads
        return batched_filter(flat_srcs, flat_refs, num_threads)

    @staticmethod
    def backward(ctx, grad_output):
        with torch.no_grad():
            srcs, refs = ctx.saved_tensors
            grad_source = grad_reference = None
            if ctx.needs_input_grad[0] and not ctx.needs_input_grad[1]:
                grad_source = batched_filter(grad_output, refs)
            if ctx.needs_input_grad[1]:
                gr, gs = [], []
                for src, ref, g in zip(srcs, refs, grad_output):
                    r, wg = _grad_reference(lattice_for(ref), src.contiguous(), ref, g.contiguous())
                    gr.append(r)
                    gs.append(wg)
                grad_reference = torch.stack(gr)
                if ctx.needs_input_grad[0]:
                    grad_source = torch.stack(gs)
        return grad_source, grad_reference, None


class LatticeGaussian(nn.Module):
    """gaussian_matrix.py:292-303: ``W @ U = filter(U) - U``."""

    def __init__(self, ref):
        super().__init__()
        self.ref = ref

    def __matmul__(self, U):
        return self(U)

    def forward(self, U):
        return LatticeFilter.apply(U, self.ref) - U


class RbfLaplacian(nn.Module):
    """gaussian_matrix.py:305-320."""

    def __init__(self, ref, normalize=True):
        super().__init__()
        self.ref = ref
        self.shape = self.ref.shape[:1] * 2
        self.normalize = normalize
        ones = torch.ones(self.shape[:1] + (1,), dtype=ref.dtype, device=ref.device)
        self.D = LatticeFilter.apply(ones, self.ref)

    def __matmul__(self, U):
        if self.normalize:
            return (U - LatticeFilter.apply(U / self.D.sqrt(), self.ref) / self.D.sqrt())
        return (self.D * U - LatticeFilter.apply(U, self.ref))


class RbfLaplacianC(LatticeGaussian):
    """gaussian_matrix.py:322-338."""

    def __init__(self, ref, normalize='sym'):
        super().__init__(ref)
        self.shape = self.ref.shape[:1] * 2
        self.normalize = normalize
        ones = torch.ones(self.shape[:1] + (1,), dtype=ref.dtype, device=ref.device)
        self.D = super().__matmul__(ones)

    def __matmul__(self, U):
        if self.normalize == 'sym':
            WsqrtDU = super().__matmul__(U / self.D.sqrt())
            return (U - WsqrtDU / self.D.sqrt())
        if self.normalize == 'right':
            return U - super().__matmul__(U / self.D)
        WU = super().__matmul__(U)
        return (self.D * U - WU)


class BatchedAdjacency(nn.Module):
    """gaussian_matrix.py:341-352: ``W(src_imgs, guide_imgs) = filter(src) - src`` on NCHW."""

    def __init__(self, num_threads=8):
        super().__init__()
        self.num_threads = num_threads

    def forward(self, src_imgs, guide_imgs):
        bs, L, h, w = src_imgs.shape
        bs, d, h, w = guide_imgs.shape
        flat_srcs = src_imgs.reshape(bs, L, -1).permute(0, 2, 1).contiguous()
        # the same guide tensor comes back on every mean-field iteration (crf_module.py:101):
        # it is flattened once (detached, cached) so its lattices are found again; when the guide
        # needs a gradient the autograd edge is added on top of the cached value
        if torch.is_grad_enabled() and guide_imgs.requires_grad:
            flat_refs = _FlattenGuide.apply(guide_imgs)
        else:
            flat_refs = flat_guide(guide_imgs)
        filtered_imgs = BatchedLatticeFilter.apply(flat_srcs, flat_refs, self.num_threads) \
            .permute(0, 2, 1).reshape(src_imgs.shape)
        return filtered_imgs - src_imgs

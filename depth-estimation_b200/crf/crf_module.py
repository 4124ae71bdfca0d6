"""Mean-field dense CRF -- mirror of crf/crf_module.py on the GPU lattice.

Same public names and signatures as the reference: ``charbonneir``,
``charbonneir2``, ``compatibility_matrix`` (:32-39), ``mean_field_infer``
(:41-53), ``potts`` (:55-64), ``charb`` (:66-79), ``CRFasRNN`` (:81-104),
``ijrgbGuide`` / ``ijGuide`` (:106-123), ``gaussian_weights(_u)`` (:8-20).

``CRFasRNN.W`` is the lattice adjacency ``BatchedAdjacency`` (the reference
imports it at :5 and it has the same ``W(src_imgs, guide_imgs)`` call shape as
the guided-filter adjacency wired at :91).  When no gradient is required the
forward runs the fused pixel-major loop of libpl_lattice (``pl_meanfield``:
softmax + compatibility in one kernel, E-update fused into slice, one lattice
per image reused by all iterations); with autograd enabled it runs the same
composition of differentiable ops as the reference.
"""
import ctypes as C

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _native as N
from .gaussian_matrix import BatchedAdjacency, lattice_for, flat_guide


def gaussian_weights(f):
    """crf_module.py:8-15 (brute-force dense kernel, tiny n only)."""
    I = torch.eye(f.shape[0]).to(f.device)
    W = torch.exp(-((f[None, :, :] - f[:, None, :]) ** 2).sum(-1)) - I
    D = W @ torch.ones(f.shape[0]).to(f.device)
    D_invsqrt = torch.diag(1 / torch.sqrt(D))
    return D_invsqrt @ W @ D_invsqrt - I


def gaussian_weights_u(f):
    """crf_module.py:17-20."""
    I = torch.eye(f.shape[0]).to(f.device)
    return torch.exp(-((f[None, :, :] - f[:, None, :]) ** 2).sum(-1)) - I


def charbonneir(a, b, gamma=.1):
    return torch.sqrt(gamma ** 2 + (a - b) ** 2) - gamma


def charbonneir2(a, b, gamma=3):
    return torch.sqrt(1 + ((a - b) / gamma) ** 2) - 1


def compatibility_matrix(compat, labels):
    return compat(labels[:, None], labels[None, :])


def mean_field_infer(E_0, W, Mu, niters=10):
    """crf_module.py:41-53: ``E = E_0 + W @ Q @ Mu`` (filter first, then Mu)."""
    Q = F.softmax(-E_0, dim=1)
    for i in range(niters):
        E = E_0 + W @ Q @ Mu
        Q = F.softmax(-E, dim=1)
    return Q


def potts(num_classes):
    """crf_module.py:55-64."""
    layer = nn.Conv2d(num_classes, num_classes, kernel_size=1, bias=False)
    with torch.no_grad():
        layer.weight.fill_(1)
        layer.weight.sub_(torch.eye(num_classes)[..., None, None])
    return layer


class charb(nn.Module):
    """crf_module.py:66-79."""

    def __init__(self, gamma):
        super().__init__()
        self.gamma = nn.Parameter(torch.tensor(gamma))
        self.s = nn.Parameter(torch.tensor(0.))

    def matrix(self, L, labels=None, device=None):
        if labels is None:
            labels = torch.arange(L, dtype=torch.float32,
                                  device=device if device is not None else self.gamma.device)
        labels = labels.reshape(-1)
        return charbonneir(labels[None, :], labels[:, None], self.gamma)

    def forward(self, x, labels=None):
        L = x.shape[1]
        Mu = self.matrix(L, labels, x.device)
        # the reference spells this F.conv2d(x, Mu[..., None, None]) (:75); cuDNN
        # convolutions run in TF32 by default, an fp32 matmul is the same contraction
        # at the reference's precision
        return torch.einsum('lk,bkhw->blhw', Mu, x) * torch.exp(self.s)

    def get_energies_from_scalar(self, x, labels):
        return charbonneir(labels, x, self.gamma * labels.max()) * torch.exp(self.s)


def _xty_accumulate(acc, a, b, chunk=4096):
    """``acc += a^T b`` for tall ``a, b [n, L]``: a batched product over row chunks keeps every SM
    busy (one [L, n] x [n, L] GEMM has L*L outputs and a 370 000-long reduction)."""
    n, L = a.shape
    k = n // chunk
    if k > 1:
        m = k * chunk
        acc += torch.bmm(a[:m].view(k, chunk, L).transpose(1, 2), b[:m].view(k, chunk, L)).sum(0)
        if m < n:
            acc.addmm_(a[m:].t(), b[m:])
    else:
        acc.addmm_(a.t(), b)


class _MeanFieldLoop(torch.autograd.Function):
    """The loop of CRFasRNN.forward (crf_module.py:97-103) as ONE differentiable op, pixel-major.

    forward : E_0 = E0;  Q_t = softmax(-E_t);  P_t = s * Q_t Mu^T;  E_{t+1} = E0 + K P_t - P_t
              (``pl_softmax_compat`` on the tensor cores + the fused filter epilogue), keeping Q_t.
    backward: exactly what autograd derives for the reference's composition -- the filter's
              gradient wrt its source is the filter itself ("Matrix is symmetric",
              gaussian_matrix.py:446), its gradient wrt the guide is gaussian_matrix.py:450-463 --
              but without the NCHW <-> [n, L] round trips and the per-op temporaries:
                  dP = K dE - dE;  dQ = s * dP Mu;  dMu += s * dP^T Q;  ds += <dP^T Q, Mu>
                  dE_prev = -Q * (dQ - rowsum(dQ * Q));  dE0 += dE
    Inputs / outputs are the module's NCHW tensors; returns E of the last iteration.
    """

    @staticmethod
    def forward(ctx, E0, refs, mu, scale, niters):
        from .gaussian_matrix import _grad_reference  # noqa: F401  (used in backward)
        B, L, H, W = E0.shape
        n, d = H * W, refs.shape[1]
        dev = E0.device
        E0c = E0.detach().float().contiguous()
        e0 = torch.empty((B, n, L), dtype=torch.float32, device=dev)
        st = N.stream_ptr(dev)
        with torch.cuda.device(dev):
            N.check(N.lib().pl_nchw_to_nlc(C.c_void_p(E0c.data_ptr()), C.c_void_p(e0.data_ptr()),
                                           B, L, n, 1.0, st))
        flat_refs = flat_guide(refs)
        mu_t = mu.detach().t().contiguous()
        s = float(scale)
        qs, ps = [], []
        keep_p = bool(ctx.needs_input_grad[1])   # the guide gradient needs the filter's source P_t again
        e_out = torch.empty_like(e0)
        for b in range(B):
            lat = lattice_for(flat_refs[b])
            e = e0[b]
            per_iter, per_iter_p = [], []
            for _ in range(niters):
                p = torch.empty((n, L), dtype=torch.float32, device=dev)
                q = torch.empty((n, L), dtype=torch.float32, device=dev)
                with torch.cuda.device(dev):
                    N.check(N.lib().pl_softmax_compat(
                        C.c_void_p(e.data_ptr()), C.c_void_p(mu_t.data_ptr()), C.c_void_p(p.data_ptr()),
                        C.c_void_p(q.data_ptr()), n, L, s, st))
                e = lat.filter(p, sub=p, add=e0[b])
                per_iter.append(q)
                if keep_p:
                    per_iter_p.append(p)
            e_out[b].copy_(e)
            qs.append(per_iter)
            ps.append(per_iter_p)
        ctx.ps = ps
        ctx.qs, ctx.flat_refs, ctx.niters, ctx.shape = qs, flat_refs, niters, (B, L, H, W, d)
        ctx.save_for_backward(mu.detach(), scale.detach() if torch.is_tensor(scale) else torch.tensor(s))
        out = torch.empty((B, L, H, W), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            N.check(N.lib().pl_nlc_to_nchw(C.c_void_p(e_out.data_ptr()), C.c_void_p(out.data_ptr()),
                                           B, L, n, 1.0, st))
        return out

    @staticmethod
    def backward(ctx, grad_out):
        from .gaussian_matrix import _grad_reference
        mu, scale = ctx.saved_tensors
        B, L, H, W, d = ctx.shape
        n = H * W
        dev = grad_out.device
        s = float(scale)
        need_e0, need_ref, need_mu, need_s = ctx.needs_input_grad[:4]
        g = grad_out.float().contiguous()
        dE_all = torch.empty((B, n, L), dtype=torch.float32, device=dev)
        st = N.stream_ptr(dev)
        with torch.cuda.device(dev):
            N.check(N.lib().pl_nchw_to_nlc(C.c_void_p(g.data_ptr()), C.c_void_p(dE_all.data_ptr()),
                                           B, L, n, 1.0, st))
        dE0 = dE_all.clone()                      # sum over t of dE_t, starting with the incoming gradient
        dref = torch.zeros((B, n, d), dtype=torch.float32, device=dev) if need_ref else None
        dM = torch.zeros((L, L), dtype=torch.float32, device=dev)     # sum_t dP_t^T Q_t (unscaled)
        mu_c = mu.contiguous()
        for b in range(B):
            lat = lattice_for(ctx.flat_refs[b])
            dE = dE_all[b]
            for t in range(ctx.niters - 1, -1, -1):
                q = ctx.qs[b][t]
                if need_ref:
                    # P_t, the filter's source: kept by the forward (pl_softmax_compat's own output)
                    p = ctx.ps[b][t] if ctx.ps[b] else torch.mm(q, mu_c.t()).mul_(s)
                    gr, wg = _grad_reference(lat, p, ctx.flat_refs[b], dE)
                    dref[b] += gr
                    dP = wg.sub_(dE)
                else:
                    dP = lat.filter(dE, sub=dE)                        # K dE - dE
                if L == 64 and dP.is_contiguous():
                    # dQ = dP Mu, dM += dP^T Q, softmax backward and dE0 += dE in one pass (pl_backward.cu)
                    with torch.cuda.device(dev):
                        N.check(N.lib().pl_meanfield_backward_step(
                            C.c_void_p(dP.data_ptr()), C.c_void_p(q.data_ptr()), C.c_void_p(mu_c.data_ptr()),
                            C.c_void_p(dE0[b].data_ptr()), C.c_void_p(dM.data_ptr()), n, L, s, st))
                    dE = dP
                    continue
                _xty_accumulate(dM, dP, q)
                dQ = torch.mm(dP, mu_c)
                # softmax(-E) backward with the scale, dE_prev = -s Q * (dQ - sum_l dQ_l Q_l), in place,
                # and dE0 += dE_prev in the same pass
                with torch.cuda.device(dev):
                    N.check(N.lib().pl_softmax_backward(
                        C.c_void_p(dQ.data_ptr()), C.c_void_p(q.data_ptr()), C.c_void_p(dQ.data_ptr()),
                        C.c_void_p(dE0[b].data_ptr()), n, L, s, st))
                dE = dQ
        grad_E0 = grad_refs = None
        if need_e0:
            grad_E0 = torch.empty((B, L, H, W), dtype=torch.float32, device=dev)
            with torch.cuda.device(dev):
                N.check(N.lib().pl_nlc_to_nchw(C.c_void_p(dE0.data_ptr()), C.c_void_p(grad_E0.data_ptr()),
                                               B, L, n, 1.0, st))
        if need_ref:
            grad_refs = dref.permute(0, 2, 1).reshape(B, d, H, W)
        grad_mu = dM * s if need_mu else None
        grad_s = (dM * mu_c).sum() if need_s else None
        return grad_E0, grad_refs, grad_mu, grad_s, None


class CRFasRNN(nn.Module):
    """crf_module.py:81-104 with the lattice pairwise term."""

    def __init__(self, mu_init, niters=5, r=20, eps=1e-5, notrain_mu=False, gaussian=False,
                 gchannels=1, fused=True):
        super().__init__()
        self.Mu = mu_init
        if notrain_mu:
            for param in self.Mu.parameters():
                param.requires_grad = False
        self.niters = niters
        # r / eps / gaussian / gchannels parametrise the reference's guided-filter
        # adjacency (:91); the lattice adjacency has no such knobs.
        self.W = BatchedAdjacency()
        self.fused = fused

    def forward(self, refs, logits, confidence=None, labels=None):
        """Assuming E0 and Refs are shape BxLxHxW and BxCxHxW"""
        if confidence is None:
            confidence = 1
        E0 = -logits * confidence
        needs_grad = torch.is_grad_enabled() and (
            E0.requires_grad or refs.requires_grad or
            any(p.requires_grad for p in self.Mu.parameters()))
        if self.fused and not needs_grad and isinstance(self.Mu, charb) and self.niters > 0:
            return self._forward_fused(refs, E0, labels)
        if self.fused and isinstance(self.Mu, charb) and self.niters > 0 and E0.is_cuda:
            # training: the same loop as one autograd op (Mu and exp(s) stay torch expressions,
            # so gamma, s, the labels and the guide's scales receive their gradients)
            L = E0.shape[1]
            mu = self.Mu.matrix(L, labels, E0.device).float()
            return -_MeanFieldLoop.apply(E0, refs, mu, torch.exp(self.Mu.s), self.niters)
        Mu = (lambda q: self.Mu(q)) if labels is None else (lambda q: self.Mu(q, labels))
        Q = F.softmax(-E0, dim=1)
        for i in range(self.niters):
            E = E0 + self.W(Mu(Q), refs)
            Q = F.softmax(-E, dim=1)
        return -E

    def depth(self, refs, logits, confidence=None, labels=None):
        """``logits2average_depth(self(refs, logits, confidence, labels), labels)``
        (mb_stereo_crf.py:62-66 on :93-104) -> ``[B, 1, H, W]``.  Without autograd the read-out
        is fused into the last slice of the mean-field loop (``pl_meanfield_depth``): the final
        logits are never written."""
        if confidence is None:
            confidence = 1
        needs_grad = torch.is_grad_enabled() and (
            logits.requires_grad or refs.requires_grad or
            any(p.requires_grad for p in self.Mu.parameters()) or
            (torch.is_tensor(confidence) and confidence.requires_grad))
        if not (self.fused and not needs_grad and isinstance(self.Mu, charb) and self.niters > 0):
            out = self.forward(refs, logits, confidence, labels)
            probs = F.softmax(out, dim=1)
            lab = (torch.arange(out.shape[1], dtype=torch.float32, device=out.device)
                   if labels is None else labels.reshape(-1).to(out.device))
            return (probs * lab[None, :, None, None]).sum(1)[:, None]
        E0 = -logits * confidence
        B, L, H, W = E0.shape
        n, d = H * W, refs.shape[1]
        dev = E0.device
        with torch.no_grad():
            mu, scale = self._compat_cached(L, labels, dev)
            flat_refs = flat_guide(refs)
            E0c = E0.float().contiguous()
            e0_nl = torch.empty((B, n, L), dtype=torch.float32, device=dev)
            with torch.cuda.device(dev):
                N.check(N.lib().pl_nchw_to_nlc(C.c_void_p(E0c.data_ptr()),
                                               C.c_void_p(e0_nl.data_ptr()), B, L, n, 1.0,
                                               N.stream_ptr(dev)))
            lab = None if labels is None else labels.reshape(-1).float().to(dev).contiguous()
            maps = [lattice_for(flat_refs[b]).mean_field_depth(e0_nl[b], mu, self.niters, scale,
                                                               labels=lab)[0] for b in range(B)]
        return torch.stack(maps).view(B, 1, H, W)

    def _compat_cached(self, L, labels, dev):
        """(Mu matrix on the device, exp(s) as a host float).  With the default labels
        (``arange(L)``) the pair is recomputed only when gamma / s change -- reading exp(s) back
        costs a host-device synchronisation per call.  With caller-supplied ``labels`` the L x L
        matrix is rebuilt on every call (it is tiny): a key on the labels tensor's address would
        match a recycled allocation holding different values (CRFdepthUpsampler builds a fresh
        ``linspace`` per frame), so only exp(s) is cached on that path."""
        key = (L, str(dev), self.Mu.gamma._version, self.Mu.s._version,
               self.Mu.gamma.data_ptr(), self.Mu.s.data_ptr())
        hit = getattr(self, "_compat_cache", None)
        if hit is None or hit[0] != key:
            mu = self.Mu.matrix(L, None, dev).float().contiguous()
            scale = float(torch.exp(self.Mu.s))
            hit = (key, mu, scale)
            self._compat_cache = hit
        if labels is not None:
            return self.Mu.matrix(L, labels.to(dev), dev).float().contiguous(), hit[2]
        return hit[1], hit[2]

    def _forward_fused(self, refs, E0, labels):
        B, L, H, W = E0.shape
        n, d = H * W, refs.shape[1]
        dev = E0.device
        with torch.no_grad():
            mu, scale = self._compat_cached(L, labels, dev)
            flat_refs = flat_guide(refs)
            E0c = E0.float().contiguous()
            e0_nl = torch.empty((B, n, L), dtype=torch.float32, device=dev)
            st = N.stream_ptr(dev)
            with torch.cuda.device(dev):
                N.check(N.lib().pl_nchw_to_nlc(C.c_void_p(E0c.data_ptr()),
                                               C.c_void_p(e0_nl.data_ptr()), B, L, n, 1.0, st))
            e_nl = torch.stack([lattice_for(flat_refs[b]).mean_field(e0_nl[b], mu, self.niters, scale)
                                for b in range(B)])
            out = torch.empty_like(E0c)
            with torch.cuda.device(dev):
                # the module returns -E (:103)
                N.check(N.lib().pl_nlc_to_nchw(C.c_void_p(e_nl.data_ptr()),
                                               C.c_void_p(out.data_ptr()), B, L, n, -1.0, st))
        return out


_IJ_CACHE = {}


def _ij_grid(bs, h, w, device):
    """``np.tile(np.mgrid[:h, :w] / np.sqrt(h**2 + w**2), (bs, 1, 1, 1))`` as fp32 on ``device``
    (crf_module.py:111-113).  Built on the device in float64 and cast, i.e. the same values the
    reference computes on the host every call, without the host work and the pageable copy;
    cached per (h, w, device)."""
    key = (h, w, str(device))
    ij = _IJ_CACHE.get(key)
    if ij is None:
        diag = float(np.sqrt(h ** 2 + w ** 2))
        ii = (torch.arange(h, dtype=torch.float64, device=device) / diag)[:, None].expand(h, w)
        jj = (torch.arange(w, dtype=torch.float64, device=device) / diag)[None, :].expand(h, w)
        ij = torch.stack([ii, jj]).float().contiguous()
        if len(_IJ_CACHE) > 8:
            _IJ_CACHE.clear()
        _IJ_CACHE[key] = ij
    return ij[None].expand(bs, 2, h, w)


class ijrgbGuide(nn.Module):
    """crf_module.py:106-114: guide = cat(ij / s_ij, rgb / s_rgb), ij = mgrid / sqrt(h^2+w^2)."""

    def __init__(self, s_ij=.1, s_rgb=.1, trainable=True):
        super().__init__()
        self.s_ij = nn.Parameter(torch.tensor(s_ij)) if trainable else s_ij
        self.s_rgb = nn.Parameter(torch.tensor(s_rgb)) if trainable else s_rgb

    def forward(self, x):
        bs, c, h, w = x.shape
        ij = _ij_grid(bs, h, w, x.device)
        return torch.cat([ij / self.s_ij, x / self.s_rgb], dim=1)


class ijGuide(nn.Module):
    """crf_module.py:116-123."""

    def __init__(self, s_ij=.1, trainable=True):
        super().__init__()
        self.s_ij = nn.Parameter(torch.tensor(s_ij)) if trainable else s_ij

    def forward(self, x):
        bs, c, h, w = x.shape
        return _ij_grid(bs, h, w, x.device) / self.s_ij


_PAIR_MU = {}
_SIDE_STREAMS = {}


def _side_stream(device, cur):
    """One side stream per (device, calling stream): callers that pipeline frames over several streams keep
    their frames independent."""
    key = (str(device), cur.cuda_stream)
    st = _SIDE_STREAMS.get(key)
    if st is None:
        if len(_SIDE_STREAMS) > 64:
            _SIDE_STREAMS.clear()
        st = torch.cuda.Stream(device)
        _SIDE_STREAMS[key] = st
    return st


def stereo_pair_depth(left, right, niters=5, gamma=.05, s=0., s_ij=.1, s_rgb=.1, window_size=9,
                      max_disp=None, criterion=None, labels=None, return_label=False):
    """Inference pipeline of the stereo notebooks with every stage on the GPU and nothing but the
    two images going in and the depth map coming out:

      cost volume   crf/depth.py:34-53 (``disparity_badness``, ``max_disp = w // 6``)  pl_cost_volume
      guide         crf_module.py:106-114 (``ijrgbGuide``), fused into the lattice build
      mean field    crf_module.py:93-104 with ``Mu = charb(gamma)``, ``exp(s)`` scale     pl_meanfield
      read-out      mb_stereo_crf.py:62-66 soft-argmax (or the arg-min label), fused into the last slice

    ``left`` / ``right``: CUDA fp32 ``[h, w, 3]`` in [0, 1] (the reference's ``read_image`` layout).
    Returns depth ``[h, w]`` (and the label map ``[h, w]`` int32 with ``return_label``).
    """
    import math
    from ..lattice import PermutohedralLattice
    from .depth import AD, cost_volume
    N.require_cuda_f32(left, "left", 3)
    h, w = int(left.shape[0]), int(left.shape[1])
    L = int(max_disp) if max_disp is not None else w // 6
    # The cost volume (all SMs, instruction bound) and the lattice build (small latency-bound kernels and one
    # host read-back) do not depend on each other: the volume goes to a side stream and overlaps the build.
    cur = torch.cuda.current_stream(left.device)
    side = _side_stream(left.device, cur)
    side.wait_stream(cur)
    with torch.cuda.stream(side):
        vol = cost_volume(left, right, L, window_size, criterion if criterion is not None else AD)
    vol.record_stream(cur)
    lab = None if labels is None else labels.reshape(-1).float().contiguous()
    if lab is None:
        # default labels 0 .. L-1: the L x L matrix depends on (L, gamma, device) only -- a streaming caller
        # would otherwise pay half a dozen tiny kernels per frame for it
        key = (L, float(gamma), str(left.device))
        mu = _PAIR_MU.get(key)
        if mu is None:
            arange = torch.arange(L, dtype=torch.float32, device=left.device)
            mu = charbonneir(arange[None, :], arange[:, None], gamma).contiguous()
            if len(_PAIR_MU) > 16:
                _PAIR_MU.clear()
            _PAIR_MU[key] = mu
    else:
        mu = charbonneir(lab[None, :], lab[:, None], gamma).contiguous()
    lat = PermutohedralLattice.from_rgb(left.permute(2, 0, 1).contiguous(), s_ij, s_rgb)
    cur.wait_stream(side)
    try:
        depth, label = lat.mean_field_depth(vol.view(h * w, L), mu, niters, math.exp(s), labels=lab,
                                            want_label=return_label)
    finally:
        lat.close()
    depth = depth.view(h, w)
    return (depth, label.view(h, w)) if return_label else depth

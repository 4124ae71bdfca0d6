"""CPU restatement of the reference's Python layer (TEST INFRASTRUCTURE, not product code).

Follows crf/gaussian_matrix.py and crf/crf_module.py line by line in torch-CPU
fp32 on top of the oracle filter (oracle/lattice_oracle.py), because the
reference modules cannot be imported as they stand (gaussian_matrix.py:8 needs
the un-vendored ``guided_filter_pytorch``; :15 JIT-loads a hard-coded path).
tests/golden/make_golden.py additionally imports the REAL reference modules
through stubs to pin this file.
"""
import numpy as np
import torch
import torch.nn.functional as F
from torch.autograd import Function

from .lattice_oracle import Oracle

_ORACLE = None


def set_backend(o):
    global _ORACLE
    _ORACLE = o


def latticefilter(src, ref):
    """gaussian_matrix.py:16 ``lattice.filter`` on CPU tensors."""
    global _ORACLE
    if _ORACLE is None:
        _ORACLE = Oracle(faithful=True)   # the reference's behaviour, growth quirk included
    out = _ORACLE.filter(src.detach().contiguous().numpy(), ref.detach().contiguous().numpy())
    return torch.from_numpy(out)


class LatticeFilter(Function):
    """gaussian_matrix.py:423-468 (without the stray print at :467)."""

    @staticmethod
    def forward(ctx, source, reference):
        assert source.shape[0] == reference.shape[0]
        ctx.save_for_backward(source, reference)
        return latticefilter(source, reference)

    @staticmethod
    def backward(ctx, grad_output):
        with torch.no_grad():
            src, ref = ctx.saved_tensors
            g = grad_output
            n, L = src.shape[-2:]
            d = ref.shape[-1]
            grad_source = grad_reference = None
            if ctx.needs_input_grad[0] and not ctx.needs_input_grad[1]:
                grad_source = latticefilter(g, ref)
            if ctx.needs_input_grad[1]:
                gf = g[..., None] * ref[..., None, :]
                sf = src[..., None] * ref[..., None, :]
                all_ = torch.cat([g, gf.reshape(n, L * d), src, sf.reshape(n, L * d)], dim=-1)
                filtered_all = latticefilter(all_, ref)
                wg, wgf, ws, wsf = torch.split(filtered_all, [L, L * d, L, L * d], dim=-1)
                grad_reference = -2 * (sf * wg[..., None] - src[..., None] * wgf.view(-1, L, d)
                                       + gf * ws[..., None] - g[..., None] * wsf.view(-1, L, d)).sum(-2)
                if ctx.needs_input_grad[0]:
                    grad_source = wg
        return grad_source, grad_reference


def batched_adjacency(src_imgs, guide_imgs):
    """gaussian_matrix.py:345-352 (BatchedAdjacency.forward), sans mp.Pool."""
    bs, L, h, w = src_imgs.shape
    d = guide_imgs.shape[1]
    flat_srcs = src_imgs.reshape(bs, L, -1).permute(0, 2, 1)
    flat_refs = guide_imgs.reshape(bs, d, -1).permute(0, 2, 1)
    filtered = torch.stack([LatticeFilter.apply(s.contiguous(), r.contiguous())
                            for s, r in zip(flat_srcs, flat_refs)])
    return filtered.permute(0, 2, 1).reshape(src_imgs.shape) - src_imgs


def charbonneir(a, b, gamma=.1):
    return torch.sqrt(gamma ** 2 + (a - b) ** 2) - gamma  # crf_module.py:32-33


def charb_forward(x, gamma, s, labels=None):
    """crf_module.py:71-75."""
    L = x.shape[1]
    if labels is None:
        labels = torch.arange(L).float()
    Mu = charbonneir(labels[None, :], labels[:, None], gamma)
    return F.conv2d(x, Mu[..., None, None]) * torch.exp(s)


def crf_as_rnn(refs, logits, gamma, s, niters, confidence=None, labels=None):
    """crf_module.py:93-104 with W = BatchedAdjacency."""
    if confidence is None:
        confidence = 1
    E0 = -logits * confidence
    Q = F.softmax(-E0, dim=1)
    for i in range(niters):
        E = E0 + batched_adjacency(charb_forward(Q, gamma, s, labels), refs)
        Q = F.softmax(-E, dim=1)
    return -E


def mean_field_infer(E_0, ref, Mu, niters=10):
    """crf_module.py:41-53 with W = LatticeGaussian(ref) (gaussian_matrix.py:292-303)."""
    Q = F.softmax(-E_0, dim=1)
    for i in range(niters):
        WQ = LatticeFilter.apply(Q, ref) - Q
        E = E_0 + WQ @ Mu
        Q = F.softmax(-E, dim=1)
    return Q


def ijrgb_guide(x, s_ij=.1, s_rgb=.1):
    """crf_module.py:111-114."""
    bs, c, h, w = x.shape
    ij = torch.from_numpy(np.tile(np.mgrid[:h, :w] / np.sqrt(h ** 2 + w ** 2), (bs, 1, 1, 1))).float()
    return torch.cat([ij / s_ij, x / s_rgb], dim=1)


def logits2average_depth(logits, labels=None):
    """mb_stereo_crf.py:62-66."""
    probs = F.softmax(logits, dim=1)
    if labels is None:
        labels = torch.arange(probs.shape[1]).float()[None, :, None, None]
    return (probs * labels).sum(1)[:, None, :, :]


# ---- stereo Module forwards (mb_stereo_crf.py:68-102,138-163), weights passed explicitly ----------

def coord_conv(x, weight, bias):
    """oil's un-vendored ``conv2d(cin, cout, coords=True)`` as tests/golden/make_golden.py and the
    product re-create it: 3x3 same-padded convolution over [x, i, j], i / j = linspace(-1, 1)."""
    b, _, h, w = x.shape
    ii = torch.linspace(-1, 1, h)[None, None, :, None].expand(b, 1, h, w)
    jj = torch.linspace(-1, 1, w)[None, None, None, :].expand(b, 1, h, w)
    return F.conv2d(torch.cat([x, ii, jj], dim=1), weight, bias, padding=1)


def crf_depth_refiner(logits, imgrgb, features, proj_w, proj_b, gamma, s, niters, confidence=None):
    """mb_stereo_crf.py:74-79 (CRFdepthRefiner.forward); with ``confidence`` :93-102."""
    guide = torch.cat((imgrgb, F.conv2d(features, proj_w, proj_b)), dim=1)
    return logits2average_depth(crf_as_rnn(guide, logits, gamma, s, niters, confidence))


def uncertainty_net(imgrgb, p):
    """mb_stereo_crf.py:84-92: coord-conv / GroupNorm(4, 16) / ReLU x 2, coord-conv -> s."""
    x = coord_conv(imgrgb, p["0__conv__weight"], p["0__conv__bias"])
    x = F.relu(F.group_norm(x, 4, p["1__weight"], p["1__bias"]))
    x = coord_conv(x, p["3__conv__weight"], p["3__conv__bias"])
    x = F.relu(F.group_norm(x, 4, p["4__weight"], p["4__bias"]))
    return coord_conv(x, p["6__conv__weight"], p["6__conv__bias"])


def crf_depth_upsampler(disp_lowres, img_highres, gamma, s, niters):
    """mb_stereo_crf.py:144-163 (CRFdepthUpsampler.forward)."""
    up = F.interpolate(disp_lowres, img_highres.size()[2:], mode='bilinear', align_corners=False)
    labels = torch.linspace(0, float(up.max()), 18).float()
    lab4 = labels[None, :, None, None]
    logits = -10 * (charbonneir(lab4, up, gamma * lab4.max()) * torch.exp(s))   # crf_module.py:77-79
    confidence = (up > 1e-2).float()
    out = crf_as_rnn(img_highres, logits, gamma, s, niters, confidence=confidence, labels=labels)
    return logits2average_depth(out, lab4)

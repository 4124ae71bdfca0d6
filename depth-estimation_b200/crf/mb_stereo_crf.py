"""Stereo CRF modules -- mirror of the nn.Module forwards in crf/mb_stereo_crf.py.

``logits2average_depth`` (:62-66), ``CRFdepthRefiner`` (:68-79),
``CRFwUncertainty`` (:81-102), ``CRFdepthUpsampler`` (:138-163).  The
``oil.Trainer`` subclasses and the ``__main__`` training script of the
reference (:14-60,105-136,165-181) are outside the hot path (they need the
un-vendored ``oil`` package) and are not reproduced.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .crf_module import CRFasRNN, charb


def logits2average_depth(logits, labels=None):
    """mb_stereo_crf.py:62-66: soft-argmax depth."""
    probs = F.softmax(logits, dim=1)
    if labels is None:
        # created on the device: the reference's host arange + .to(device) is a pageable H2D
        # copy, which blocks the calling thread until the stream has drained
        labels = torch.arange(probs.shape[1], dtype=torch.float32,
                              device=probs.device)[None, :, None, None]
    average_depths = (probs * labels).sum(1)
    return average_depths[:, None, :, :]


class CRFdepthRefiner(nn.Module):
    """mb_stereo_crf.py:68-79."""

    def __init__(self, d_in=64, d_guide=16, r=15, niters=2, eps=1e-2, gamma=.05):
        super().__init__()
        self.CRF = CRFasRNN(charb(gamma), niters=niters, r=r, eps=eps, gchannels=d_guide)
        self.projection = nn.Conv2d(d_in, d_guide - 3, kernel_size=1)

    def forward(self, inputs):
        logits, imgrgb, features = inputs
        projected_features = self.projection(features)
        guide = torch.cat((imgrgb, projected_features), dim=1)
        return self.CRF.depth(guide, logits)      # = logits2average_depth(self.CRF(guide, logits))


class _CoordConv2d(nn.Module):
    """The reference's ``conv2d(cin, cout, coords=True)`` comes from the
    un-vendored ``oil`` package (mb_stereo_crf.py:5-7): a 3x3 same-padded
    convolution that appends normalised (i, j) coordinate channels."""

    def __init__(self, cin, cout):
        super().__init__()
        self.conv = nn.Conv2d(cin + 2, cout, kernel_size=3, padding=1)

    def forward(self, x):
        b, _, h, w = x.shape
        ii = torch.linspace(-1, 1, h, device=x.device)[None, None, :, None].expand(b, 1, h, w)
        jj = torch.linspace(-1, 1, w, device=x.device)[None, None, None, :].expand(b, 1, h, w)
        return self.conv(torch.cat([x, ii, jj], dim=1))


class CRFwUncertainty(CRFdepthRefiner):
    """mb_stereo_crf.py:81-102: per-pixel confidence = exp(-s)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.uncertainty_net = nn.Sequential(
            _CoordConv2d(3, 16), nn.GroupNorm(4, 16), nn.ReLU(),
            _CoordConv2d(16, 16), nn.GroupNorm(4, 16), nn.ReLU(),
            _CoordConv2d(16, 1),
        )

    def forward(self, inputs):
        logits, imgrgb, features = inputs
        projected_features = self.projection(features)
        guide = torch.cat((imgrgb, projected_features), dim=1)
        s = self.uncertainty_net(imgrgb)
        confidence = torch.exp(-s)
        depth_out = self.CRF.depth(guide, logits, confidence)
        return depth_out, confidence


class CRFdepthUpsampler(nn.Module):
    """mb_stereo_crf.py:138-163: 18 linspace labels, logits = -10 * charbonnier energies."""

    def __init__(self, d_in=64, d_guide=3, r=15, niters=2, eps=1e-2, gamma=.05):
        super().__init__()
        self.CRF = CRFasRNN(charb(gamma), niters=niters, r=r, eps=eps, gchannels=d_guide)

    def forward(self, inputs):
        disp_lowres, img_highres, vgg_features = inputs
        upsampled_disp = F.interpolate(disp_lowres, img_highres.size()[2:], mode='bilinear',
                                       align_corners=False)
        guide = img_highres
        max_depth = upsampled_disp.max()
        labels = torch.linspace(0, float(max_depth), 18).float().to(upsampled_disp.device)
        logits = -10 * self.CRF.Mu.get_energies_from_scalar(upsampled_disp,
                                                            labels[None, :, None, None])
        confidence = (upsampled_disp > 1e-2).float()
        return self.CRF.depth(guide, logits, confidence=confidence, labels=labels)

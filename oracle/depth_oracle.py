"""TEST INFRASTRUCTURE -- CPU restatement of crf/depth.py:18-53 (float64 numpy, like the
reference).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.

Pinned by tests/golden/cost_volume_*.npz, produced by the reference's own
``crf.depth.disparity_badness`` (tests/golden/make_golden.py section 7).
"""
import numpy as np


def SD(a, b):          # depth.py:18-19
    return (a - b) ** 2


def AD(a, b):          # depth.py:20-21
    return np.abs(a - b)


def nprod(a, b):       # depth.py:22-23
    return -1 * a * b


def _reflect(i, n):
    """scipy.ndimage 'reflect' boundary (d c b a | a b c d | d c b a)."""
    i = np.asarray(i)
    period = 2 * n
    i = np.mod(i, period)
    return np.where(i >= n, period - 1 - i, i)


def disparity_badness(img1, img2, window_size=9, criterion=AD, max_disp=None):
    """depth.py:34-53.  ``max_disp`` defaults to the reference's ``w // 6`` (:40)."""
    ws = window_size
    h, w, c = img1.shape
    if max_disp is None:
        max_disp = w // 6
    padded = np.pad(img2, ((0, 0), (max_disp, 0), (0, 0)), mode='constant')      # :44
    out = np.zeros((h, w, max_disp))
    for i in range(max_disp):                                                     # :47-49
        shifted = padded[:, max_disp - i:w + max_disp - i]
        out[:, :, i] = criterion(img1, shifted).sum(2)
    # :50-51  scipy.ndimage.convolve(out, ones((ws, ws, 1))), default mode='reflect', origin 0:
    # for an odd window the box is centred, offsets -ws//2 .. ws//2
    if ws % 2 == 0:
        raise ValueError("odd windows only (the reference's default is 9)")
    r = ws // 2
    lo, hi = -r, r
    ys = _reflect(np.arange(h)[:, None] + np.arange(lo, hi + 1)[None, :], h)     # [h, ws]
    xs = _reflect(np.arange(w)[:, None] + np.arange(lo, hi + 1)[None, :], w)     # [w, ws]
    agg = out[ys].sum(1)               # vertical window:   [h, ws, w, L] -> [h, w, L]
    agg = agg[:, xs].sum(2)            # horizontal window: [h, w, ws, L] -> [h, w, L]
    return agg


def disparity_estimate(img1, img2, window_size=9, criterion=AD):
    return np.argmin(disparity_badness(img1, img2, window_size, criterion), axis=-1)

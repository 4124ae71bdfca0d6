"""Seeded synthetic inputs of the BASELINE.json shapes.

Images are smooth-plus-texture RGB (sum of low-frequency cosines + 5 % noise) so
the lattice occupancy M/n lands in the range real photographs give rather than
the i.i.d. worst case; unaries are |N(0,1)|*4 with a piecewise-constant true
label at zero cost.  numpy only: usable by the tests, the bench and the oracle.
"""
import numpy as np


def image(h, w, seed=0, noise=0.05):
    """``noise`` = amplitude of the uniform per-pixel texture (0.05).  With the sharp
    bilateral scales 0.05 gives M/n ~ 0.05 at 1 MP and 0.4 gives M/n ~ 0.4 (heavily textured)."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[:h, :w].astype(np.float32)
    img = np.zeros((h, w, 3), np.float32)
    for c in range(3):
        for _ in range(6):
            fx, fy = rng.uniform(0.5, 4, 2)
            ph = rng.uniform(0, 2 * np.pi, 2)
            a = rng.uniform(0.1, 0.3)
            img[..., c] += (a * np.cos(2 * np.pi * fx * xx / w + ph[0])
                            * np.cos(2 * np.pi * fy * yy / h + ph[1])).astype(np.float32)
    img -= img.min()
    img /= img.max()
    img += np.float32(noise) * rng.random(img.shape).astype(np.float32)
    return np.clip(img, 0, 1).astype(np.float32)


def ijrgb_features(img, s_ij=0.1, s_rgb=0.1, rgb_first=False):
    """[h*w, 5] features: crf_module.py:111-114 order (i, j, r, g, b), or the
    (r, g, b, i, j) order of gaussian_matrix.py:500-505 when rgb_first."""
    h, w, _ = img.shape
    ij = (np.mgrid[:h, :w] / np.sqrt(h ** 2 + w ** 2)).transpose(1, 2, 0)
    parts = [img / s_rgb, ij / s_ij] if rgb_first else [ij / s_ij, img / s_rgb]
    return np.concatenate(parts, -1).reshape(h * w, 5).astype(np.float32)


def disparity(h, w, L, seed=0):
    """Piecewise-constant integer disparity in [0, L)."""
    rng = np.random.default_rng(seed + 1)
    coarse = rng.integers(0, L, size=((h + 63) // 64, (w + 63) // 64))
    return np.kron(coarse, np.ones((64, 64), dtype=np.int64))[:h, :w]


def unary(h, w, L, seed=0):
    """E0[h*w, L]: |N(0,1)|*4, true label's cost 0."""
    rng = np.random.default_rng(seed + 2)
    e = np.abs(rng.standard_normal((h * w, L)).astype(np.float32)) * 4
    e[np.arange(h * w), disparity(h, w, L, seed).reshape(-1)] = 0
    return e


def charb_matrix(L, gamma=0.05):
    """crf_module.py:32-33,74 with labels = arange(L)."""
    lab = np.arange(L, dtype=np.float32)
    return (np.sqrt(np.float32(gamma) ** 2 + (lab[None, :] - lab[:, None]) ** 2)
            - np.float32(gamma)).astype(np.float32)

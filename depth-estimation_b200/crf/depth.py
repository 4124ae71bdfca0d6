"""Plane-sweep matching costs -- mirror of crf/depth.py:18-53, on the GPU.

``disparity_badness(img1, img2, window_size=9, criterion=AD)`` has the reference's signature and
returns the same ``[h, w, max_disp]`` volume (``max_disp = w // 6``, depth.py:40); the work is one
launch of ``pl_cost_volume`` (csrc/pl_cost.cu) instead of a numpy loop over disparities plus
``scipy.ndimage.convolve``.  numpy images in -> numpy volume out (staged through the GPU, fp32
arithmetic, returned as float64 like the reference); CUDA tensors in -> CUDA fp32 tensor out,
already in the pixel-major layout the mean-field loop consumes (``volume.view(h * w, L)``).

``AD``, ``SD`` and ``nprod`` are the reference's criteria (depth.py:18-23); they are callable
like the originals and double as selectors for the fused kernel.
"""
import ctypes as C

import numpy as np
import torch

from .. import _native as N


class _Criterion:
    def __init__(self, code, fn, name):
        self.code, self.fn, self.__name__ = code, fn, name

    def __call__(self, imga, imgb):
        return self.fn(imga, imgb)


SD = _Criterion(1, lambda a, b: (a - b) ** 2, "SD")              # depth.py:18-19
AD = _Criterion(0, lambda a, b: abs(a - b), "AD")               # depth.py:20-21
nprod = _Criterion(2, lambda a, b: -1 * a * b, "nprod")      # depth.py:22-23


def cost_volume(left, right, max_disp, window_size=9, criterion=AD):
    """``left``, ``right``: CUDA fp32 ``[h, w, c]`` -> CUDA fp32 ``[h, w, max_disp]``."""
    N.require_cuda_f32(left, "left", 3)
    N.require_cuda_f32(right, "right", 3)
    if left.shape != right.shape:
        raise ValueError(f"image shapes differ: {tuple(left.shape)} vs {tuple(right.shape)}")
    if not isinstance(criterion, _Criterion):
        raise TypeError("criterion must be AD, SD or nprod from this module")
    left, right = left.contiguous(), right.contiguous()
    h, w, c = (int(v) for v in left.shape)
    out = torch.empty((h, w, int(max_disp)), dtype=torch.float32, device=left.device)
    with torch.cuda.device(left.device):
        N.check(N.lib().pl_cost_volume(C.c_void_p(left.data_ptr()), C.c_void_p(right.data_ptr()),
                                       C.c_void_p(out.data_ptr()), h, w, c, int(max_disp),
                                       int(window_size), criterion.code, N.stream_ptr(left.device)))
    return out


def disparity_badness(img1, img2, window_size=9, criterion=AD):
    """depth.py:34-53: the energy of each disparity 0 .. w//6 - 1 from the window sweep."""
    is_np = isinstance(img1, np.ndarray)
    if is_np:
        dev = torch.device("cuda", torch.cuda.current_device())
        a = torch.from_numpy(np.ascontiguousarray(img1, dtype=np.float32)).to(dev)
        b = torch.from_numpy(np.ascontiguousarray(img2, dtype=np.float32)).to(dev)
    else:
        a, b = img1.float(), img2.float()
    max_disp = int(a.shape[1]) // 6
    out = cost_volume(a, b, max_disp, window_size, criterion)
    return out.cpu().numpy().astype(np.float64) if is_np else out


def disparity_estimate(img1, img2, window_size=9, criterion=AD):
    """depth.py:29-32: minimum-energy disparity of the window sweep."""
    energy = disparity_badness(img1, img2, window_size, criterion)
    if isinstance(energy, np.ndarray):
        return np.argmin(energy, axis=-1)
    return torch.argmin(energy, dim=-1)

"""The two small-lattice blur forms of csrc/pl_filter.cu -- blur_flow_kernel (no barrier between the
passes: threads poll per-pass scratch buffers pre-filled with a sentinel) and blur_reg_kernel (one
grid barrier per pass) -- perform the same fp32 operations in the same order (permutohedral.h:503-522),
so the filter must come out BIT-identical.  PL_BLUR_FORM=reg (read per call) selects the barrier form.

The splat in front of the blur accumulates with float atomics, whose order is not repeatable; the
sources here are sparse enough (one non-zero pixel in 997) that a vertex sees at most two addends
from zero, which commute exactly -- asserted by requiring each form to repeat itself bit for bit.

Consecutive calls on one lattice with 64, 16, 64, 4 and 32 channels exercise the sentinel bookkeeping
of the two scratch sets (a call refills the set the previous call dirtied, over the previous call's
extent).
"""
import importlib
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _features(h, w, s_ij=0.1, s_rgb=0.1, noise=0.05):
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    return synthetic.ijrgb_features(synthetic.image(h, w, 3, noise), s_ij, s_rgb)


class _Form:
    def __init__(self, name):
        self.name, self.old = name, None

    def __enter__(self):
        self.old = os.environ.get("PL_BLUR_FORM")
        if self.name is None:
            os.environ.pop("PL_BLUR_FORM", None)
        else:
            os.environ["PL_BLUR_FORM"] = self.name

    def __exit__(self, *exc):
        if self.old is None:
            os.environ.pop("PL_BLUR_FORM", None)
        else:
            os.environ["PL_BLUR_FORM"] = self.old


def _sparse(n, C, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    src = torch.zeros((n, C), device="cuda")
    rows = torch.arange(seed % 7, n, 997, device="cuda")
    src[rows] = torch.rand((rows.numel(), C), generator=g, device="cuda") + 0.5
    return src


@pytest.mark.parametrize("h,w,kw", [(120, 160, {}), (32, 48, {"s_ij": 0.02, "s_rgb": 0.2, "noise": 0.3}),
                                     # M/n ~ 3.7: 64 channels take the flow form with two items per thread,
                                     # fewer channels the one-item form
                                     (48, 64, {"s_ij": 0.02, "s_rgb": 0.2, "noise": 0.3})])
def test_flow_and_barrier_blur_are_bit_identical(native, h, w, kw):
    feats = torch.from_numpy(_features(h, w, **kw)).cuda()
    lat = native.PermutohedralLattice(feats)
    try:
        n = h * w
        assert lat.num_vertices * 4 <= 148 * 1024, "shape no longer takes the small-lattice forms"
        for i, C in enumerate([64, 16, 64, 4, 32]):
            src = _sparse(n, C, 11 + i)
            with _Form("reg"):
                a = lat.filter(src).clone()
                a2 = lat.filter(src).clone()
            with _Form(None):
                b = lat.filter(src).clone()
                b2 = lat.filter(src).clone()
            assert torch.equal(a, a2) and torch.equal(b, b2), "splat order leaked into the result: source not sparse enough"
            assert torch.isfinite(b).all()
            assert int((b != 0).sum()) > n * C // 50          # the blur has spread the sparse source
            assert torch.equal(a, b), f"C = {C}: max |diff| {(a - b).abs().max().item():.3g}"
    finally:
        lat.close()


def test_mean_field_loop_rotation_matches_barrier_form(native):
    """5 mean-field iterations: with the flow form the loop rotates three value buffers (the blur zeroes
    the next splat target inside the kernel); with the barrier form two.  Dense inputs, so the splat's
    atomics make the two runs differ by rounding only -- same bound as two runs of one form."""
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    h, w, L = 120, 160, 64
    feats = torch.from_numpy(_features(h, w)).cuda()
    g = torch.Generator(device="cuda").manual_seed(5)
    e0 = torch.randn((h * w, L), generator=g, device="cuda").abs_()
    mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).cuda()
    lat = native.PermutohedralLattice(feats)
    try:
        with _Form("reg"):
            a = lat.mean_field(e0, mu, 5, 0.01).clone()
            a2 = lat.mean_field(e0, mu, 5, 0.01).clone()
        with _Form(None):
            b = lat.mean_field(e0, mu, 5, 0.01).clone()
            b2 = lat.mean_field(e0, mu, 5, 0.01).clone()
        scale = a.abs().max().item()
        noise = max((a - a2).abs().max().item(), (b - b2).abs().max().item())
        assert torch.isfinite(b).all()
        # same bound as the loop parity tests (2e-4 of the scale) unless the run-to-run spread is larger still
        assert (a - b).abs().max().item() <= max(4 * noise, 2e-4 * scale)
        assert (a.argmin(1) == b.argmin(1)).float().mean().item() > 0.9995
    finally:
        lat.close()


def test_streaming_frames_do_not_grow_device_memory(native):
    """A new lattice per frame, closed after its mean-field call: free device memory must level off (the
    third value buffer of the rotating loop once leaked: 2.2 MB per frame, and every ~15 frames the pool
    went back to the driver for more, stalling the stream for 0.3 - 100 ms)."""
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    h, w, L = 120, 160, 64
    feats = torch.from_numpy(_features(h, w)).cuda()
    g = torch.Generator(device="cuda").manual_seed(5)
    e0 = torch.randn((h * w, L), generator=g, device="cuda").abs_()
    mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).cuda()

    def frame():
        lat = native.PermutohedralLattice(feats)
        lat.mean_field(e0, mu, 3, 0.01)
        lat.filter(e0[:, :16].contiguous())
        lat.close()

    for _ in range(20):
        frame()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(200):
        frame()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free0 - free1 < (8 << 20), f"device memory grew by {(free0 - free1) >> 20} MB over 200 frames"

"""pl_softmax_compat_noalloc (include/pl_lattice.h): the allocation-free variant of pl_softmax_compat
whose scratch is a static __device__ array inside the library.  Called straight through the C ABI and
compared with the existing entry point on the same inputs, on consecutive calls that leave different
contents in the shared scratch (L = 64, then 32, then 64 again, then 16).

Collected last on purpose; passes alone in a fresh process and after
the other tests (it depends on no state but the library's own static array).
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-6   # of the output scale: the tolerance tests/test_crf_gpu.py holds this transform to


def _call(fn, e, mu_t, scale, want_q):
    n, L = e.shape
    p = torch.empty_like(e)
    q = torch.empty_like(e) if want_q else None
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = fn(C.c_void_p(e.data_ptr()), C.c_void_p(mu_t.data_ptr()), C.c_void_p(p.data_ptr()),
            C.c_void_p(q.data_ptr() if want_q else 0), n, L, C.c_float(scale), st)
    return rc, p, q


def test_noalloc_matches_existing_entry_point_on_consecutive_calls(native):
    lib = native._native.lib()
    g = torch.Generator(device="cuda").manual_seed(7)
    for i, (n, L, scale, want_q) in enumerate([(5000, 64, 1.3, True), (3001, 32, 0.7, False),
                                                (5000, 64, 1.0, False), (777, 16, 2.0, True)]):
        e = torch.randn((n, L), generator=g, device="cuda") * 3
        lab = torch.arange(L, dtype=torch.float32, device="cuda") * (1.0 + 0.25 * i)
        mu_t = (torch.sqrt(0.05 ** 2 + (lab[None, :] - lab[:, None]) ** 2) - 0.05).t().contiguous()
        rc0, p0, q0 = _call(lib.pl_softmax_compat, e, mu_t, scale, want_q)
        assert rc0 == 0, lib.pl_last_error().decode()
        for rep in range(2):     # two calls in a row on the same scratch
            rc1, p1, q1 = _call(lib.pl_softmax_compat_noalloc, e, mu_t, scale, want_q)
            assert rc1 == 0, lib.pl_last_error().decode()
            torch.cuda.synchronize()
            ref = p0.abs().max().item()
            assert (p1 - p0).abs().max().item() <= TOL * ref, (n, L, rep)
            if want_q:
                assert (q1 - q0).abs().max().item() <= TOL
        # against the definition as well: P = scale * softmax(-E) Mu^T in float64
        want = (torch.softmax(-e.double(), dim=1) @ mu_t.double()) * scale
        assert (p1.double() - want).abs().max().item() <= 1e-5 * want.abs().max().item()


def test_noalloc_rejects_sizes_beyond_the_static_scratch(native):
    lib = native._native.lib()
    e = torch.zeros((256, 128), device="cuda")
    mu_t = torch.zeros((128, 128), device="cuda")
    rc, _, _ = _call(lib.pl_softmax_compat_noalloc, e, mu_t, 1.0, False)
    assert rc == 1                                     # PL_ERR_INVALID
    msg = lib.pl_last_error().decode()
    assert "exceeds the static scratch" in msg and "64" in msg
    e = torch.zeros((256, 24), device="cuda")          # within the limit but not a supported L
    mu_t = torch.zeros((24, 24), device="cuda")
    rc, _, _ = _call(lib.pl_softmax_compat_noalloc, e, mu_t, 1.0, False)
    assert rc == 1

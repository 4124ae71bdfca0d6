"""Large label counts (L = 128, 256) and label sharding on the GPU.

* pl_softmax_compat_blocked (pl_compat_tc.cu): tcgen05 3xTF32 contraction with K = L, N = shard
  width, blocked (all-gathered) E layout -- against fp64.
* the label-sharded mean-field loop, G ranks emulated one after another in one process on one
  GPU (ranks must not spin on each other on a single device), against the unsharded loop and
  the CPU oracle.
"""
import ctypes as C
import importlib

import numpy as np
import pytest
import torch

synthetic = importlib.import_module("depth_estimation_b200.synthetic")
pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("L,G", [(128, 1), (128, 2), (256, 1), (256, 4), (256, 8), (64, 2), (192, 3)])
def test_blocked_compat_tensor_core(native, L, G):
    N = native._native
    w = L // G
    for n in (1, 130, 777):
        g = torch.Generator().manual_seed(L * 7 + G + n)
        e = torch.randn((n, L), generator=g) * 3
        mu = torch.from_numpy(synthetic.charb_matrix(L)) + torch.rand((L, L), generator=g) * 0.1
        q = torch.softmax(-e.double(), dim=1)
        want = (q @ mu.double().t()) * 0.9
        blocks = e.reshape(n, G, w).permute(1, 0, 2).contiguous().cuda()     # [G][n][w]
        mt = mu.t().contiguous().cuda()
        for rank in range(G):
            p = torch.full((n, w), float("nan"), device="cuda")
            N.check(N.lib().pl_softmax_compat_blocked(
                C.c_void_p(blocks.data_ptr()), G, C.c_void_p(mt.data_ptr()), C.c_void_p(p.data_ptr()),
                n, L, rank * w, w, 0.9, N.stream_ptr(p.device)))
            torch.cuda.synchronize()
            assert _rel(p, want[:, rank * w:(rank + 1) * w]) < 1e-5, (L, G, n, rank)   # fast exp + online softmax rescaling
    assert N.lib().pl_debug_tc_timeouts() == 0


@pytest.mark.parametrize("n,L", [(60001, 128), (40000, 256), (45000, 192)])
def test_compat_many_tiles_per_cta(native, n, L):
    """More tiles than CTAs: the double-buffered operand planes / accumulators, the next-tile row maxima
    and the Q output rescale all wrap around several times; ragged last tile."""
    N = native._native
    g = torch.Generator(device="cuda").manual_seed(n + L)
    e = torch.randn((n, L), generator=g, device="cuda") * 4
    e[::7] += 30.0          # rows far from zero: the row maximum matters
    e[3::11, : L // 2] -= 25.0
    mu = (torch.from_numpy(synthetic.charb_matrix(L)) + torch.rand((L, L)) * 0.1).cuda()
    want_q = torch.softmax(-e.double(), dim=1)
    want_p = (want_q @ mu.double().t()) * 0.8
    mt = mu.t().contiguous()
    p, q = torch.full_like(e, float("nan")), torch.full_like(e, float("nan"))
    N.check(N.lib().pl_softmax_compat(C.c_void_p(e.data_ptr()), C.c_void_p(mt.data_ptr()),
                                      C.c_void_p(p.data_ptr()), C.c_void_p(q.data_ptr()),
                                      n, L, 0.8, N.stream_ptr(e.device)))
    torch.cuda.synchronize()
    assert _rel(q, want_q) < 1e-5
    assert _rel(p, want_p) < 1e-5
    # per-row check as well: a wrong row sum or maximum on a few rows would hide in the global scale
    row_err = ((p.double() - want_p).abs().amax(1) / want_p.abs().amax(1)).max()
    assert float(row_err) < 2e-5
    assert N.lib().pl_debug_tc_timeouts() == 0


class _FakeGroupRun:
    """Runs LabelShardedMeanField's iteration for all ranks in lock step on one device."""

    def __init__(self, native, lat, world):
        ls = importlib.import_module("depth_estimation_b200.crf.label_shards")
        self.ls, self.world = ls, world
        self.backend = ls.CudaBackend(lat)

    def run(self, e0, mu, niters, scale):
        n, L = e0.shape
        w = L // self.world
        mu_t = mu.t().contiguous()
        shards0 = [e0[:, r * w:(r + 1) * w].contiguous() for r in range(self.world)]
        shards = list(shards0)
        for _ in range(niters):
            blocks = torch.stack(shards)                       # what the all-gather delivers
            shards = [self.backend.filter_update(
                self.backend.compat(blocks, mu_t, r * w, w, scale), shards0[r])
                for r in range(self.world)]
        return torch.cat(shards, dim=1)


@pytest.mark.parametrize("h,w,L,G,niters", [(24, 32, 128, 4, 3), (20, 28, 256, 8, 2), (20, 28, 256, 2, 2)])
def test_label_sharded_mean_field_matches_unsharded_and_oracle(native, oracle, h, w, L, G, niters):
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=L))
    e0 = synthetic.unary(h, w, L, seed=L)
    mu = synthetic.charb_matrix(L) * 0.3
    hdl = oracle.build(ref)
    e = e0.copy()
    for _ in range(niters):
        x = -e - (-e).max(1, keepdims=True)
        q = np.exp(x) / np.exp(x).sum(1, keepdims=True)
        p = (q @ mu.T).astype(np.float32) * np.float32(1.1)
        e = e0 + (hdl.apply(p) - p)
    lat = native.PermutohedralLattice(torch.from_numpy(ref).cuda())
    e0d, mud = torch.from_numpy(e0).cuda(), torch.from_numpy(mu).cuda()
    unsharded = lat.mean_field(e0d, mud, niters, 1.1).cpu().numpy()          # large-L path
    cuda_core = lat.mean_field(e0d, mud, niters, 1.1, no_tensor=True).cpu().numpy()
    sharded = _FakeGroupRun(native, lat, G).run(e0d, mud, niters, 1.1).cpu().numpy()
    scale = np.abs(e).max()
    for got in (unsharded, cuda_core, sharded):
        assert np.abs(got - e).max() / scale < 2e-4
        assert (got.argmin(1) == e.argmin(1)).mean() > 0.999
    assert native._native.lib().pl_debug_tc_timeouts() == 0


@pytest.mark.parametrize("L,G", [(128, 2), (256, 4), (192, 3)])
def test_compat_with_separate_block_pointers(native, L, G):
    """pl_softmax_compat_peers: the G label blocks live in G separate allocations (what peer
    mappings look like to the kernel) -- same result as the contiguous blocked layout."""
    N = native._native
    w, n = L // G, 515
    g = torch.Generator().manual_seed(L + G)
    e = torch.randn((n, L), generator=g) * 3
    mu = torch.from_numpy(synthetic.charb_matrix(L)) + torch.rand((L, L), generator=g) * 0.1
    want = (torch.softmax(-e.double(), dim=1) @ mu.double().t()) * 1.1
    # separate tensors, deliberately allocated in reverse order with gaps in between
    parts, pad = [None] * G, []
    for b in reversed(range(G)):
        pad.append(torch.empty(1000 + 37 * b, device="cuda"))
        parts[b] = e[:, b * w:(b + 1) * w].contiguous().cuda()
    mt = mu.t().contiguous().cuda()
    ptrs = (C.c_void_p * G)(*[t.data_ptr() for t in parts])
    for rank in range(G):
        p = torch.full((n, w), float("nan"), device="cuda")
        N.check(N.lib().pl_softmax_compat_peers(ptrs, G, C.c_void_p(mt.data_ptr()), C.c_void_p(p.data_ptr()),
                                                n, L, rank * w, w, 1.1, N.stream_ptr(p.device)))
        torch.cuda.synchronize()
        assert _rel(p, want[:, rank * w:(rank + 1) * w]) < 1e-5, (L, G, rank)
    assert N.lib().pl_debug_tc_timeouts() == 0


def test_peer_memory_backend_single_process(native):
    """world = 1: the IPC buffers, the torch views over them and the iteration order (the
    multi-process form is tools/peer_check.py under torchrun, run by test_peer_shards_two_gpus)."""
    ls = importlib.import_module("depth_estimation_b200.crf.label_shards")
    h, w, L, niters = 40, 56, 128, 3
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=1), 0.1, 0.1)
    e0 = torch.from_numpy(synthetic.unary(h, w, L, seed=1)).cuda()
    mu = torch.from_numpy(synthetic.charb_matrix(L) * 0.7).cuda()
    lat = native.PermutohedralLattice(torch.from_numpy(ref).cuda())
    want = lat.mean_field(e0, mu, niters, 1.1)
    backend = ls.PeerMemoryBackend(lat, h * w, L)
    try:
        for _ in range(2):       # twice: the buffer parity carries over between runs
            got = backend.run(e0, mu, niters, 1.1).clone()
            assert _rel(got, want) < 2e-5
    finally:
        backend.close()


@pytest.mark.parametrize("L,G", [(128, 2), (256, 4), (256, 8), (64, 1)])
def test_compat_shards_emulated_on_one_gpu(native, L, G):
    """pl_softmax_compat_shards with the G label blocks of E and P all on this GPU: each emulated rank
    processes its pixel shard and writes the P rows into every owner's block; the assembled P equals
    the plain transform (crf_module.py:99-102 with :71-75)."""
    import ctypes as C
    ls = importlib.import_module("depth_estimation_b200.crf.label_shards")
    N = native._native
    n, w = 1000, L // G
    g = torch.Generator(device="cuda").manual_seed(L + G)
    e = torch.randn((n, L), generator=g, device="cuda") * 3
    mu = torch.from_numpy(synthetic.charb_matrix(L) * 0.6).cuda()
    mu_t = mu.t().contiguous()
    want = (torch.softmax(-e.double(), 1) @ mu_t.double() * 1.2).float()
    eb = [e[:, b * w:(b + 1) * w].contiguous() for b in range(G)]
    pb = [torch.full((n, w), float("nan"), device="cuda") for _ in range(G)]
    ep = (C.c_void_p * G)(*[t.data_ptr() for t in eb])
    pp = (C.c_void_p * G)(*[t.data_ptr() for t in pb])
    for rank in range(G):
        row0, cnt = ls.pixel_shard(n, G, rank)
        N.check(N.lib().pl_softmax_compat_shards(ep, pp, G, C.c_void_p(mu_t.data_ptr()), row0, cnt, L,
                                                 C.c_float(1.2), N.stream_ptr(e.device)))
    torch.cuda.synchronize()
    got = torch.cat(pb, dim=1)
    assert torch.isfinite(got).all()
    assert _rel(got, want) < 1e-5
    assert N.lib().pl_debug_tc_timeouts() == 0


def test_peer_transpose_single_process(native):
    """world = 1 PeerTransposeMeanField: IPC buffers, fences and iteration order against the plain loop."""
    ls = importlib.import_module("depth_estimation_b200.crf.label_shards")
    h, w, L, niters = 40, 56, 128, 3
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=1), 0.1, 0.1)
    e0 = torch.from_numpy(synthetic.unary(h, w, L, seed=1)).cuda()
    mu = torch.from_numpy(synthetic.charb_matrix(L) * 0.7).cuda()
    lat = native.PermutohedralLattice(torch.from_numpy(ref).cuda())
    want = lat.mean_field(e0, mu, niters, 1.1)
    pt = ls.PeerTransposeMeanField(h * w, L, e0.device)
    try:
        for _ in range(2):
            assert _rel(pt.run(lat, e0, mu, niters, 1.1), want) < 2e-5
        tr = ls.TransposeExchangeMeanField(ls.CudaBackend(lat)).run(e0, mu, niters, 1.1)
        assert _rel(tr, want) < 2e-5
    finally:
        pt.close()


def test_peer_shards_all_gpus(native):
    """One process per GPU on every GPU of the box (2, 4 or 8): NCCL all-gather, peer-memory reads and
    both transpose-exchange forms against the unsharded loop (skipped only on a single-GPU box)."""
    import os
    import subprocess
    import sys
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs two GPUs")
    world = 8 if ngpu >= 8 else (4 if ngpu >= 4 else 2)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", "29731", os.path.join(root, "tools", "peer_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "peer_check ok" in out.stdout

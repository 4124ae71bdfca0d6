"""Host-side logic of the multi-GPU partitionings on CPU: world_size-2 and -4 gloo groups, with the
CPU oracle standing in for the per-rank compute (tests may use the oracle; the product back end
is CUDA-only).  Checks partition arithmetic, the gathered block layout and the iteration order
of LabelShardedMeanField against the unsharded oracle loop.
"""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

synthetic = importlib.import_module("depth_estimation_b200.synthetic")
ls = importlib.import_module("depth_estimation_b200.crf.label_shards")


class OracleBackend:
    def __init__(self, ref):
        from oracle.lattice_oracle import Oracle
        self.h = Oracle(faithful=False).build(ref)

    def compat(self, e_blocks, mu_t, l0, count, scale):
        world, n, w = e_blocks.shape
        e = e_blocks.permute(1, 0, 2).reshape(n, world * w)          # label = block * w + offset
        q = torch.softmax(-e, dim=1)
        return (q @ mu_t[:, l0:l0 + count]) * scale                   # sum_k Mu[l][k] q[k]

    def compat_rows(self, e_blocks, mu_t, scale):
        world, rows, w = e_blocks.shape
        e = e_blocks.permute(1, 0, 2).reshape(rows, world * w)
        return (torch.softmax(-e, dim=1) @ mu_t) * scale

    def filter_update(self, p, e0):
        f = torch.from_numpy(self.h.apply(np.ascontiguousarray(p.numpy())))
        return e0 + (f - p)


def _unsharded(ref, e0, mu, niters, scale):
    b = OracleBackend(ref)
    e = torch.from_numpy(e0)
    mu_t = torch.from_numpy(mu).t().contiguous()
    for _ in range(niters):
        p = (torch.softmax(-e, dim=1) @ mu_t) * scale
        e = b.filter_update(p, torch.from_numpy(e0))
    return e.numpy()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, h, w, L, niters, out_dir, mode="gather"):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=3))
        e0 = synthetic.unary(h, w, L, seed=3)
        mu = synthetic.charb_matrix(L) * 0.5
        l0, cnt = ls.label_shard(L, world, rank)
        cls = ls.LabelShardedMeanField if mode == "gather" else ls.TransposeExchangeMeanField
        mf = cls(OracleBackend(ref))
        e = mf.run(torch.from_numpy(np.ascontiguousarray(e0[:, l0:l0 + cnt])), torch.from_numpy(mu),
                   niters, 1.2)
        np.save(os.path.join(out_dir, f"shard{rank}.npy"), e.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_label_sharded_mean_field_gloo(tmp_path, world):
    h, w, L, niters = 16, 20, 8, 3
    mp.spawn(_worker, args=(world, _free_port(), h, w, L, niters, str(tmp_path)), nprocs=world,
             join=True)
    got = np.concatenate([np.load(tmp_path / f"shard{r}.npy") for r in range(world)], axis=1)
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=3))
    want = _unsharded(ref, synthetic.unary(h, w, L, seed=3), synthetic.charb_matrix(L) * 0.5, niters, 1.2)
    assert np.abs(got - want).max() / np.abs(want).max() < 1e-5


@pytest.mark.parametrize("world,h,w", [(2, 16, 20), (4, 16, 20), (4, 9, 14)])
def test_transpose_exchange_mean_field_gloo(tmp_path, world, h, w):
    """Pixel shards for softmax + compat, label shards for the filter, two all-to-alls per iteration
    (h * w = 126 leaves the last ranks with a short / empty pixel shard)."""
    L, niters = 8, 3
    mp.spawn(_worker, args=(world, _free_port(), h, w, L, niters, str(tmp_path), "transpose"),
             nprocs=world, join=True)
    got = np.concatenate([np.load(tmp_path / f"shard{r}.npy") for r in range(world)], axis=1)
    ref = synthetic.ijrgb_features(synthetic.image(h, w, seed=3))
    want = _unsharded(ref, synthetic.unary(h, w, L, seed=3), synthetic.charb_matrix(L) * 0.5, niters, 1.2)
    assert np.abs(got - want).max() / np.abs(want).max() < 1e-5


def test_pixel_shard_arithmetic():
    for n, world in [(1000, 4), (126, 4), (2900000, 8), (128, 2), (1, 3)]:
        rows = [ls.pixel_shard(n, world, r) for r in range(world)]
        assert rows[0][0] == 0 and sum(c for _, c in rows) == n
        for (a, ca), (b, _) in zip(rows, rows[1:]):
            assert a + ca == b and (ca % 128 == 0 or b == n)


def test_partition_arithmetic():
    for batch, world in [(64, 8), (7, 4), (3, 8), (1, 1)]:
        seen = []
        for r in range(world):
            sl = ls.image_shard(batch, world, r)
            seen += list(range(batch))[sl]
        assert seen == list(range(batch))
    assert ls.label_shard(256, 8, 3) == (96, 32)
    with pytest.raises(ValueError):
        ls.label_shard(18, 4, 0)

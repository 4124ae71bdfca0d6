"""Multi-GPU partitioning of the mean-field loop (one process per GPU, torch.distributed).

The reference's only parallelism is a process pool over the images of a batch
(crf/gaussian_matrix.py:370-377).  Two partitionings exist here:

* image-parallel -- ``image_shard``: independent stereo pairs per rank, no collective on the
  data path (what ``bench.py --gpus N`` runs).
* label-channel sharding -- ``LabelShardedMeanField``: for large disparity counts every rank
  builds the SAME lattice from the same guide and owns ``L / G`` consecutive labels of E0 / E.
  The softmax and the compatibility transform couple all labels, so each iteration all-gathers
  the E shards (NCCL over NVLink; one collective per iteration, n * L floats received per rank),
  every rank recomputes the softmax over all L labels and contracts it with ITS rows of Mu, then
  filters its own channels on its own copy of the lattice:
      E_g = E0_g + (filter(P_g) - P_g),   P_g = scale * softmax(-E) Mu[shard_g, :]^T
  (crf/crf_module.py:100-102 restricted to the rank's label rows).  Pixel sharding is not
  offered: the lattice couples the whole image (replicas only).

The loop is written against a small compute back end so the host-side logic (partition,
gather layout, iteration order) is testable on CPU with gloo; the default back end is the CUDA
library.
"""
import ctypes as C

import torch
import torch.distributed as dist

from .. import _native as N


def image_shard(batch, world, rank):
    """Contiguous slice of a batch of ``batch`` images for ``rank`` (image-parallel)."""
    per, extra = divmod(batch, world)
    start = rank * per + min(rank, extra)
    return slice(start, start + per + (1 if rank < extra else 0))


def label_shard(L, world, rank):
    """(first label, count) owned by ``rank``; L must divide evenly."""
    if L % world != 0:
        raise ValueError(f"{L} labels do not split evenly over {world} ranks")
    return rank * (L // world), L // world


class CudaBackend:
    """Per-rank compute on the GPU library (pl_softmax_compat_blocked + cached lattice)."""

    def __init__(self, lattice):
        self.lat = lattice

    def compat(self, e_blocks, mu_t, l0, count, scale):
        world, n, w = e_blocks.shape
        L = world * w
        p = torch.empty((n, count), dtype=torch.float32, device=e_blocks.device)
        with torch.cuda.device(e_blocks.device):
            N.check(N.lib().pl_softmax_compat_blocked(
                C.c_void_p(e_blocks.data_ptr()), world, C.c_void_p(mu_t.data_ptr()),
                C.c_void_p(p.data_ptr()), n, L, l0, count, float(scale),
                N.stream_ptr(e_blocks.device)))
        return p

    def filter_update(self, p, e0):
        """E0 + (filter(P) - P) on this rank's channels."""
        return self.lat.filter(p, sub=p, add=e0)


class LabelShardedMeanField:
    """``run(e0_shard[n, L/G], mu[L, L], niters, scale) -> E_shard[n, L/G]`` (last iteration's E,
    the module returns -E).  ``group`` is a torch.distributed process group (None = default)."""

    def __init__(self, backend, group=None):
        self.backend = backend
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def gather(self, e_shard):
        """[n, w] per rank -> blocked [G, n, w] on every rank (rank g's shard in block g)."""
        e_shard = e_shard.contiguous()
        out = torch.empty((self.world,) + tuple(e_shard.shape), dtype=e_shard.dtype,
                          device=e_shard.device)
        if self.world == 1:
            out[0].copy_(e_shard)
        else:
            # flat views: block g of the output is rank g's shard (gloo and nccl agree on this)
            dist.all_gather_into_tensor(out.view(-1), e_shard.view(-1), group=self.group)
        return out

    def run(self, e0_shard, mu, niters, scale=1.0):
        n, w = e0_shard.shape
        L = w * self.world
        if tuple(mu.shape) != (L, L):
            raise ValueError(f"mu must be [{L}, {L}] for {self.world} shards of {w} labels")
        l0, count = label_shard(L, self.world, self.rank)
        mu_t = mu.t().contiguous()
        e = e0_shard
        for _ in range(int(niters)):
            blocks = self.gather(e)
            p = self.backend.compat(blocks, mu_t, l0, count, scale)
            e = self.backend.filter_update(p, e0_shard)
        return e

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


class _RawDeviceBuffer:
    """A cudaMalloc'ed block exposed to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, shape):
        self.ptr, self.shape = int(ptr), tuple(shape)
        self.__cuda_array_interface__ = {"shape": self.shape, "typestr": "<f4", "data": (self.ptr, False),
                                         "version": 2, "strides": None}


class PeerMemoryBackend:
    """Label shards exchanged through PEER MEMORY instead of an all-gather (one process per GPU on
    one node, NVLink / NVSwitch): every rank keeps its E shard in a CUDA-IPC buffer that the other
    ranks map; ``pl_softmax_compat_peers`` then reads all G shards in place -- the remote ones over
    NVLink -- inside the kernel that computes the softmax statistics and the compatibility GEMM.
    Nothing is gathered into a contiguous copy (the all-gather writes n*L floats to HBM and the
    kernel reads them back); the only collective left is a 4-byte all-reduce per iteration that
    orders the ranks on the stream (shards complete before anybody reads them).

    Two shard buffers alternate between iterations so a rank may start writing E_{t+1} while a
    slower peer still reads E_t.
    """

    def __init__(self, lattice, n, width, group=None):
        self.lat, self.group = lattice, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n, self.w = int(n), int(width)
        self.dev = lattice.device
        nbytes = self.n * self.w * 4
        self._own, self._views, handles = [], [], []
        with torch.cuda.device(self.dev):
            for _ in range(2):
                ptr, h = C.c_void_p(), C.create_string_buffer(64)
                N.check(N.lib().pl_peer_alloc(nbytes, C.byref(ptr), h))
                self._own.append(ptr.value)
                handles.append(bytes(h.raw))
                self._views.append(torch.as_tensor(_RawDeviceBuffer(ptr.value, (self.n, self.w)),
                                                   device=self.dev))
        everyone = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(everyone, handles, group=group)
        else:
            everyone[0] = handles
        self._opened = []
        self._ptrs = []          # per buffer parity: ctypes array of G block pointers
        with torch.cuda.device(self.dev):
            for k in range(2):
                arr = (C.c_void_p * self.world)()
                for g in range(self.world):
                    if g == self.rank:
                        arr[g] = self._own[k]
                    else:
                        ptr = C.c_void_p()
                        N.check(N.lib().pl_peer_open(everyone[g][k], C.byref(ptr)))
                        self._opened.append(ptr.value)
                        arr[g] = ptr.value
                self._ptrs.append(arr)
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self._turn = 0

    def shard_buffer(self):
        """The buffer this rank's next E shard goes into (a torch view of the IPC block)."""
        return self._views[self._turn]

    def publish(self):
        """All ranks' shards of the current turn are complete after this returns (in stream order)."""
        if self.world > 1:
            dist.all_reduce(self._flag, group=self.group)

    def compat(self, mu_t, l0, count, scale):
        L = self.world * self.w
        p = torch.empty((self.n, count), dtype=torch.float32, device=self.dev)
        with torch.cuda.device(self.dev):
            N.check(N.lib().pl_softmax_compat_peers(
                self._ptrs[self._turn], self.world, C.c_void_p(mu_t.data_ptr()),
                C.c_void_p(p.data_ptr()), self.n, L, l0, count, float(scale), N.stream_ptr(self.dev)))
        self._turn ^= 1
        return p

    def run(self, e0_shard, mu, niters, scale=1.0):
        """Same contract as ``LabelShardedMeanField.run``."""
        n, w = e0_shard.shape
        if (n, w) != (self.n, self.w):
            raise ValueError("shard shape differs from the one the buffers were made for")
        L = w * self.world
        l0, count = label_shard(L, self.world, self.rank)
        mu_t = mu.t().contiguous()
        self.shard_buffer().copy_(e0_shard)
        e = e0_shard
        for _ in range(int(niters)):
            self.publish()
            p = self.compat(mu_t, l0, count, scale)
            # E_g = E0_g + K P_g - P_g written straight into the buffer the peers will read
            e = self.lat.filter(p, sub=p, add=e0_shard, out=self.shard_buffer())
        return e

    def close(self):
        with torch.cuda.device(self.dev):
            torch.cuda.synchronize(self.dev)
            if self.world > 1:
                dist.barrier(group=self.group)      # nobody unmaps while a peer may still read
            for ptr in self._opened:
                N.lib().pl_peer_close(C.c_void_p(ptr))
            if self.world > 1:
                dist.barrier(group=self.group)
            self._views = []
            for ptr in self._own:
                N.lib().pl_peer_free(C.c_void_p(ptr))
        self._opened, self._own = [], []


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

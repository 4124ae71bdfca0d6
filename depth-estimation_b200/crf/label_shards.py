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


# ---- pixel / label transpose exchange ----------------------------------------------------------------
# Softmax and the compatibility transform couple the LABELS of one pixel; the lattice filter couples
# the PIXELS of one label channel.  So instead of all-gathering the whole energy volume on every rank
# (n * L floats received per rank and iteration, G-fold redundant softmax), shard pixels for softmax +
# compat and labels for the filter, and transpose between the two with two all-to-alls per iteration:
#   E label shard [n, L/G]  --all-to-all-->  E pixel shard, blocked [G, n/G, L/G]
#   P pixel shard [n/G, L]  --all-to-all-->  P label shard [n, L/G]
# Per rank that is 2 * n * L * 4 * (G - 1) / G^2 bytes each way (0.65 GB at 2.9 MP x 256, G = 8, against
# 2.6 GB for the all-gather), and nobody repeats anybody's softmax.

def pixel_shard(n, world, rank, align=128):
    """(first row, count) of the pixel rows ``rank`` owns for softmax + compat: contiguous, sized in
    multiples of the 128-pixel tile; the last ranks may get fewer (or no) rows."""
    per = -(-n // world)
    per = -(-per // align) * align
    row0 = min(n, rank * per)
    return row0, min(n, row0 + per) - row0


class TransposeExchangeMeanField:
    """``run(e0_shard[n, L/G], mu[L, L], niters, scale) -> E_shard[n, L/G]`` with the transpose
    exchange over ``torch.distributed`` all-to-alls (NCCL over NVLink; emulated with all-gathers where
    the back end has no all-to-all, i.e. gloo on CPU for the host-logic tests)."""

    def __init__(self, backend, group=None):
        self.backend, self.group = backend, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def _all_to_all(self, out, inp, out_splits, in_splits):
        if self.world == 1:
            out.copy_(inp)
            return
        if dist.get_backend(self.group) == "nccl":
            dist.all_to_all_single(out, inp, out_splits, in_splits, group=self.group)
            return
        # gloo: every rank publishes its whole send buffer (padded to a common size), receivers cut out
        # their pieces -- same data movement as seen by the caller, for the CPU tests only
        sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(self.world)]
        dist.all_gather(sizes, torch.tensor([inp.numel()], dtype=torch.int64), group=self.group)
        cap = max(1, int(max(int(s) for s in sizes)))
        padded = torch.zeros(cap, dtype=inp.dtype)
        padded[: inp.numel()] = inp
        bufs = [torch.zeros(cap, dtype=inp.dtype) for _ in range(self.world)]
        all_splits = [torch.zeros(self.world, dtype=torch.int64) for _ in range(self.world)]
        dist.all_gather(bufs, padded, group=self.group)
        dist.all_gather(all_splits, torch.tensor(in_splits, dtype=torch.int64), group=self.group)
        pos = 0
        for src in range(self.world):
            off = int(all_splits[src][: self.rank].sum())
            cnt = int(all_splits[src][self.rank])
            assert cnt == out_splits[src]
            out[pos: pos + cnt] = bufs[src][off: off + cnt]
            pos += cnt

    def run(self, e0_shard, mu, niters, scale=1.0):
        n, w = e0_shard.shape
        G = self.world
        L = w * G
        if tuple(mu.shape) != (L, L):
            raise ValueError(f"mu must be [{L}, {L}] for {G} shards of {w} labels")
        rows = [pixel_shard(n, G, r) for r in range(G)]
        mine = rows[self.rank][1]
        mu_t = mu.t().contiguous()
        e = e0_shard.contiguous()
        for _ in range(int(niters)):
            # E: label shard -> this rank's pixel rows of every label block
            blocks = torch.empty((G, mine, w), dtype=e.dtype, device=e.device)
            self._all_to_all(blocks.view(-1), e.view(-1), [mine * w] * G, [c * w for _, c in rows])
            p_loc = self.backend.compat_rows(blocks, mu_t, scale)            # [mine, L]
            # P: pixel shard, cut by destination label block -> label shard for all pixels
            packed = p_loc.view(mine, G, w).permute(1, 0, 2).contiguous()
            p = torch.empty((n, w), dtype=e.dtype, device=e.device)
            self._all_to_all(p.view(-1), packed.view(-1), [c * w for _, c in rows], [mine * w] * G)
            e = self.backend.filter_update(p, e0_shard)
        return e


def _compat_rows_cuda(self, e_blocks, mu_t, scale):
    """All L output labels for the pixel rows of a blocked E [G, rows, L/G] (pl_softmax_compat_blocked)."""
    world, rows, w = e_blocks.shape
    L = world * w
    p = torch.empty((rows, L), dtype=torch.float32, device=e_blocks.device)
    if rows > 0:
        with torch.cuda.device(e_blocks.device):
            N.check(N.lib().pl_softmax_compat_blocked(
                C.c_void_p(e_blocks.data_ptr()), world, C.c_void_p(mu_t.data_ptr()),
                C.c_void_p(p.data_ptr()), rows, L, 0, L, float(scale), N.stream_ptr(e_blocks.device)))
    return p


CudaBackend.compat_rows = _compat_rows_cuda


class PeerTransposeMeanField:
    """The same exchange with NO collective on the data path: every rank keeps its label block of E and
    of P in CUDA-IPC buffers that the peers map, and ONE kernel per iteration
    (``pl_softmax_compat_shards``) reads this rank's pixel rows of every E block in place -- the remote
    ones over NVLink -- runs softmax + the compatibility GEMM on tcgen05, and writes each P tile
    straight into its label owner's buffer.  Two 4-byte NCCL all-reduces per iteration order the ranks
    on the stream (E complete before the kernel, P complete after)."""

    def __init__(self, n, width, device, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n, self.w, self.dev = int(n), int(width), device
        nbytes = max(16, self.n * self.w * 4)
        self._own, self._views, handles = [], [], []
        with torch.cuda.device(self.dev):
            for _ in range(2):                       # [0] E block, [1] P block
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
        self._opened, self._ptrs = [], []
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

    def _fence(self):
        if self.world > 1:
            dist.all_reduce(self._flag, group=self.group)

    def run(self, lat, e0_shard, mu, niters, scale=1.0):
        n, w = e0_shard.shape
        if (n, w) != (self.n, self.w):
            raise ValueError("shard shape differs from the one the buffers were made for")
        L = w * self.world
        row0, cnt = pixel_shard(n, self.world, self.rank)
        mu_t = mu.t().contiguous()
        e_buf, p_buf = self._views
        e_buf.copy_(e0_shard)
        for _ in range(int(niters)):
            self._fence()                                # every rank's E block is complete
            with torch.cuda.device(self.dev):
                N.check(N.lib().pl_softmax_compat_shards(
                    self._ptrs[0], self._ptrs[1], self.world, C.c_void_p(mu_t.data_ptr()), row0, cnt, L,
                    float(scale), N.stream_ptr(self.dev)))
            self._fence()                                # every rank's P block is complete
            lat.filter(p_buf, sub=p_buf, add=e0_shard, out=e_buf)
        return e_buf.clone()

    def close(self):
        with torch.cuda.device(self.dev):
            torch.cuda.synchronize(self.dev)
            if self.world > 1:
                dist.barrier(group=self.group)
            for ptr in self._opened:
                N.lib().pl_peer_close(C.c_void_p(ptr))
            if self.world > 1:
                dist.barrier(group=self.group)
            self._views = []
            for ptr in self._own:
                N.lib().pl_peer_free(C.c_void_p(ptr))
        self._opened, self._own = [], []

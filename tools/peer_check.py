"""torchrun --nproc-per-node G tools/peer_check.py : label sharding on G GPUs -- NCCL all-gather,
peer-memory reads, the pixel / label transpose exchange over NCCL all-to-alls and its one-kernel
peer-memory form -- against the unsharded single-GPU loop (same inputs on every rank)."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import depth_estimation_b200 as pkg

ls = importlib.import_module("depth_estimation_b200.crf.label_shards")
synthetic = importlib.import_module("depth_estimation_b200.synthetic")


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    h, w, L, niters = 96, 128, min(64 * world, 256), 3   # the transpose forms produce all L labels per kernel: L <= 256
    ref = torch.from_numpy(synthetic.ijrgb_features(synthetic.image(h, w, seed=2), 0.1, 0.1)).to(dev)
    e0 = torch.from_numpy(synthetic.unary(h, w, L, seed=2)).to(dev)
    mu = torch.from_numpy(synthetic.charb_matrix(L) * 0.7).to(dev)
    lat = pkg.PermutohedralLattice(ref)
    want = lat.mean_field(e0, mu, niters, 1.1)
    l0, cnt = ls.label_shard(L, world, rank)
    shard0 = e0[:, l0:l0 + cnt].contiguous()
    gathered = ls.LabelShardedMeanField(ls.CudaBackend(lat)).run(shard0, mu, niters, 1.1)
    backend = ls.PeerMemoryBackend(lat, h * w, cnt)
    errs = []
    for _ in range(2):
        got = backend.run(shard0, mu, niters, 1.1).clone()
        ref_shard = want[:, l0:l0 + cnt]
        scale = float(want.abs().max())
        errs.append(float((got - ref_shard).abs().max()) / scale)
        errs.append(float((got - gathered).abs().max()) / scale)
    backend.close()
    # pixel / label transpose exchange: NCCL all-to-alls, then the one-kernel peer-memory form
    tr = ls.TransposeExchangeMeanField(ls.CudaBackend(lat)).run(shard0, mu, niters, 1.1)
    errs.append(float((tr - want[:, l0:l0 + cnt]).abs().max()) / float(want.abs().max()))
    pt = ls.PeerTransposeMeanField(h * w, cnt, dev)
    for _ in range(2):
        got = pt.run(lat, shard0, mu, niters, 1.1)
        errs.append(float((got - want[:, l0:l0 + cnt]).abs().max()) / float(want.abs().max()))
    pt.close()
    t = torch.tensor([max(errs)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    timeouts = pkg._native.lib().pl_debug_tc_timeouts()
    ok = float(t) < 5e-5 and timeouts == 0
    if rank == 0:
        print(f"peer_check {'ok' if ok else 'FAILED'}: world {world}, max rel err {float(t):.3e}, "
              f"tc timeouts {timeouts}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

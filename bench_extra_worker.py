"""Child process of `bench.py --gpus N` (N > 1): runs bench_extra.multi_gpu_extras in its own process group.

The label-sharded modes read and write peer-GPU memory inside persistent tensor-core kernels whose mbarrier
waits trap on time-out.  A trap poisons the CUDA context, so these extras run in a child process per rank:
if one fails, bench.py still prints its JSON line (headline, e2e, roofline) with the failure recorded in
`extra.configs_error`.  Rank 0 writes the entries to the file given as argv[1].
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import depth_estimation_b200 as pkg
    import bench_extra as X
    out = X.multi_gpu_extras(pkg, dev, rank, local, world)
    torch.cuda.synchronize(dev)
    if rank == 0:
        with open(sys.argv[1], "w") as f:
            json.dump(out, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

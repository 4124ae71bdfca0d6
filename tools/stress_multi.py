"""torchrun --nproc-per-node G tools/stress_multi.py [reps]: repeat bench_extra.multi_gpu_extras (label-sharded
config 4 in the four exchange forms + image-parallel config 5) to shake out intermittent failures."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import depth_estimation_b200 as pkg
import bench_extra as X
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
for i in range(reps):
    t0 = time.time()
    out = X.multi_gpu_extras(pkg, dev, rank, local, world, steps=2, warmup=1)
    torch.cuda.synchronize(dev)
    if rank == 0:
        bad = [c["exchange"] for c in out if "unavailable" in c]
        print(f"[stress] rep {i} ok in {time.time() - t0:.1f} s, unavailable: {bad}", file=sys.stderr, flush=True)
dist.barrier()
dist.destroy_process_group()

"""Where the lattice build spends its time at large n (PL_BUILD_TRACE=1 prints the phases)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench_configs as bc
import depth_estimation_b200 as pkg
dev = torch.device("cuda", 0)
for (h, w) in [(2048, 2048), (4096, 2048)]:
    feats = bc._features(h, w, 1, dev)
    for rep in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        lat = pkg.PermutohedralLattice(feats)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        lat.close()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        print(f"{h}x{w} rep {rep}: build {1e3*(t1-t0):.3f} ms, close {1e3*(t2-t1):.3f} ms", file=sys.stderr)

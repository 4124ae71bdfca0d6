"""One workload with the fused read-out, a few frames (for ncu): python tools/r2_depth_one.py H W L variant [frames]"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import depth_estimation_b200 as pkg
import bench_util as U
synthetic = importlib.import_module("depth_estimation_b200.synthetic")
h, w, L, variant = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
frames = int(sys.argv[5]) if len(sys.argv) > 5 else 3
dev = torch.device("cuda", 0)
feats = torch.from_numpy(U.features(h, w, variant)).to(dev)
g = torch.Generator(device=dev).manual_seed(1)
e0 = torch.randn((h * w, L), generator=g, device=dev).abs_().mul_(4)
mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).to(dev)
for _ in range(frames):
    lat = pkg.PermutohedralLattice(feats)
    out = lat.mean_field_depth(e0, mu, 5, 1.0, labels=torch.arange(L, dtype=torch.float32, device=dev))
    M = lat.num_vertices
    lat.close()
torch.cuda.synchronize()
print("ok", M)
if len(sys.argv) > 6:      # timing: frames with the read-out against frames without, lattice cached
    lat = pkg.PermutohedralLattice(feats)
    lab = torch.arange(L, dtype=torch.float32, device=dev)
    def run(fn, reps=60):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) * 1000 / reps
    for rep in range(2):
        t4 = run(lambda: lat.mean_field_depth(e0, mu, 4, 1.0, labels=lab))
        t5 = run(lambda: lat.mean_field_depth(e0, mu, 5, 1.0, labels=lab))
        print("rep", rep, "5 iterations + read-out %.1f us; one iteration %.1f us" % (t5, t5 - t4), flush=True)
    lat.close()

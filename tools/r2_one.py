"""One workload, a few frames (for ncu): python tools/r2_one.py H W L variant [frames]"""
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
    out = lat.mean_field(e0, mu, 5, 1.0)
    M = lat.num_vertices
    lat.close()
torch.cuda.synchronize()
print("ok", M)

"""Host-side time of the three calls of a frame (constructor incl. the wait for M, mean_field, close);
prints the distribution and the frames where mean_field or close took unusually long."""
import importlib, os, sys, time, gc
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import depth_estimation_b200 as pkg
import bench_util as U
synthetic = importlib.import_module("depth_estimation_b200.synthetic")
h, w, L = 500, 740, 64
dev = torch.device("cuda", 0)
feats = torch.from_numpy(U.features(h, w, "default")).to(dev)
g = torch.Generator(device=dev).manual_seed(1)
e0 = torch.randn((h * w, L), generator=g, device=dev).abs_().mul_(4)
mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).to(dev)
if len(sys.argv) > 1 and sys.argv[1] == "nogc":
    gc.disable()
rows = []
pc = time.perf_counter
for i in range(640):
    t0 = pc(); lat = pkg.PermutohedralLattice(feats)
    t1 = pc(); out = lat.mean_field(e0, mu, 5, 1.0)
    t2 = pc(); lat.close()
    t3 = pc()
    if i >= 40:
        rows.append((i, (t1 - t0) * 1e6, (t2 - t1) * 1e6, (t3 - t2) * 1e6))
torch.cuda.synchronize()
import statistics
for k, name in ((1, "ctor"), (2, "mean_field"), (3, "close")):
    v = sorted(r[k] for r in rows)
    print("%-10s median %.0f  p90 %.0f  p99 %.0f  max %.0f us" % (name, v[len(v) // 2], v[int(len(v) * .9)], v[int(len(v) * .99)], v[-1]))
slow = [r for r in rows if r[2] > 150 or r[3] > 100]
print("slow frames (index, ctor, mean_field, close):", [(r[0], int(r[1]), int(r[2]), int(r[3])) for r in slow[:40]])

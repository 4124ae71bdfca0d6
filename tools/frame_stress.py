"""Spread of the whole-frame time (new lattice per frame) over batches of 100 frames in ONE process:
python tools/frame_stress.py [batches]"""
import importlib, os, sys
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

def frame():
    lat = pkg.PermutohedralLattice(feats)
    out = lat.mean_field(e0, mu, 5, 1.0)
    lat.close()
    return out

for _ in range(5):
    frame()
out = []
for b in range(int(sys.argv[1]) if len(sys.argv) > 1 else 12):
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(101)]
    evs[0].record()
    for i in range(100):
        frame()
        evs[i + 1].record()
    torch.cuda.synchronize()
    t = [evs[i].elapsed_time(evs[i + 1]) * 1000 for i in range(100)]
    t.sort()
    out.append("%.0f/%.0f/%.0f" % (t[50], t[95], t[99]))
print("median/p95/max per batch (us):", " ".join(out), flush=True)

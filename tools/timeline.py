"""Device timeline of a few whole frames (kernels, memsets, copies: start, duration, gap to the previous
one) through torch.profiler / CUPTI:  python tools/timeline.py [frames] [cached]"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
import depth_estimation_b200 as pkg
import bench_util as U
synthetic = importlib.import_module("depth_estimation_b200.synthetic")
h, w, L = 500, 740, 64
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 3
cached_mode = len(sys.argv) > 2 and sys.argv[2] == "cached"
dev = torch.device("cuda", 0)
feats = torch.from_numpy(U.features(h, w, "default")).to(dev)
g = torch.Generator(device=dev).manual_seed(1)
e0 = torch.randn((h * w, L), generator=g, device=dev).abs_().mul_(4)
mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).to(dev)
cached = pkg.PermutohedralLattice(feats)

def frame():
    if cached_mode:
        return cached.mean_field(e0, mu, 5, 1.0)
    lat = pkg.PermutohedralLattice(feats)
    out = lat.mean_field(e0, mu, 5, 1.0)
    lat.close()
    return out

for _ in range(10):
    frame()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(frames):
        frame()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
prev_end = t0
for e in ev:
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    print("%9.1f us  +%6.1f gap  %7.1f us  %s" % (s, e.time_range.start - prev_end, d, e.name[:70]))
    prev_end = max(prev_end, e.time_range.end)
print("total %.1f us for %d frames" % (prev_end - t0, frames))

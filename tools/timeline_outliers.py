"""CUPTI timeline of many whole frames; prints only the long device activities and the long gaps:
python tools/timeline_outliers.py [frames] [threshold_us]"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
import depth_estimation_b200 as pkg
import bench_util as U
synthetic = importlib.import_module("depth_estimation_b200.synthetic")
h, w, L = 500, 740, 64
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 300
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 400.0
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

for _ in range(10):
    frame()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(frames):
        frame()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
prev_end, prev_name = t0, ""
n_long = 0
for e in ev:
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    gap = e.time_range.start - prev_end
    if d > thr or gap > thr:
        n_long += 1
        if n_long <= 40:
            print("%10.0f us  gap %8.0f  dur %8.0f  %s   (after %s)" % (s, gap, d, e.name[:48], prev_name[:32]))
    prev_end = max(prev_end, e.time_range.end)
    prev_name = e.name
print("frames %d, total %.0f us = %.0f us per frame, %d long activities / gaps" % (frames, prev_end - t0, (prev_end - t0) / frames, n_long))

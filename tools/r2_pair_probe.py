"""Where the stereo_pair_depth e2e time goes (scratch driver)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import depth_estimation_b200 as pkg
cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
syn = importlib.import_module("depth_estimation_b200.synthetic")
N = pkg._native
dev = torch.device("cuda", 0)
img = syn.image(500, 740, 0)
left = torch.from_numpy(img).to(dev); right = torch.from_numpy(np.roll(img, -5, 1).copy()).to(dev)
f = lambda: cm.stereo_pair_depth(left, right, niters=5, gamma=0.05, max_disp=64)
for _ in range(5): f()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); a.record()
for _ in range(50): f()
b.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
print("device-resident pair path: %.3f ms/frame by events, %.3f ms wall" % (a.elapsed_time(b) / 50, (t1 - t0) * 20))
t0 = time.perf_counter()
for _ in range(50): f()
t_enq = (time.perf_counter() - t0) * 20
torch.cuda.synchronize()
print("host enqueue time per frame: %.3f ms" % t_enq)
N.profile(True); f(); torch.cuda.synchronize(); N.profile(False)
rep = N.profile_report()
print({k: round(v[1], 4) for k, v in sorted(rep.items(), key=lambda kv: -kv[1][1])}, "sum %.3f" % sum(v[1] for v in rep.values()))

import importlib, sys, time, os
sys.path.insert(0, os.getcwd())
import torch
import bench
import depth_estimation_b200 as pkg
cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
mb = importlib.import_module("depth_estimation_b200.crf.mb_stereo_crf")
gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
dev = torch.device("cuda", 0)
H, W, L, D = bench.H, bench.W, bench.L, bench.D
img, feats, e0, mu = bench._inputs(0)
crf = cm.CRFasRNN(cm.charb(bench.GAMMA), niters=5).to(dev)
guide_h = torch.from_numpy(feats).reshape(H, W, D).permute(2, 0, 1)[None].contiguous().pin_memory()
logits_h = (-torch.from_numpy(e0)).reshape(H, W, L).permute(2, 0, 1)[None].contiguous().pin_memory()
g = guide_h.to(dev); lg = logits_h.to(dev)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
def h2d():
    a = guide_h.to(dev, non_blocking=True); b = logits_h.to(dev, non_blocking=True)
print("h2d ms", t(h2d))
def comp():
    with torch.no_grad():
        d = mb.logits2average_depth(crf(g, lg))
    torch.cuda.synchronize(); gm.clear_lattice_cache()
print("module compute ms", t(comp))
def crf_only():
    with torch.no_grad():
        d = crf(g, lg)
    torch.cuda.synchronize(); gm.clear_lattice_cache()
print("crf only ms", t(crf_only))
with torch.no_grad():
    out = crf(g, lg)
def readout():
    with torch.no_grad():
        d = mb.logits2average_depth(out)
print("readout ms", t(readout))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    comp()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=60))

import importlib, sys, time, os, cProfile, pstats
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
g = torch.from_numpy(feats).reshape(H, W, D).permute(2, 0, 1)[None].contiguous().to(dev)
lg = (-torch.from_numpy(e0)).reshape(H, W, L).permute(2, 0, 1)[None].contiguous().to(dev)
def frame():
    with torch.no_grad():
        d = mb.logits2average_depth(crf(g, lg))
    torch.cuda.synchronize()
    gm.forget(g)
for _ in range(5): frame()
pr = cProfile.Profile(); pr.enable()
for _ in range(50): frame()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)

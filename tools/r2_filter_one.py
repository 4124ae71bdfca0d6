"""One filter apply (for ncu): python tools/r2_filter_one.py H W C variant"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import depth_estimation_b200 as pkg
import bench_util as U
h, w, C, v = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
f = torch.from_numpy(U.features(h, w, v)).cuda()
lat = pkg.PermutohedralLattice(f)
st = lat.export()
vid = st["vid"].long()
T = vid.shape[0] // 128
x = vid[:T * 128].reshape(T, 768).sort(dim=1).values
Ud = ((x[:, 1:] != x[:, :-1]).sum(1) + 1).float()
print("M", lat.num_vertices, "U mean %.1f p50 %.0f p99 %.0f max %.0f" % (Ud.mean(), Ud.median(), torch.quantile(Ud, 0.99), Ud.max()))
src = torch.rand((h * w, C), device="cuda")
out = torch.empty_like(src)
for _ in range(3):
    lat.filter(src, out=out)
torch.cuda.synchronize()

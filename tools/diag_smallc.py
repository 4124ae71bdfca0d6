"""Tiled vs flat splat/slice kernels at small channel counts (1 MP, cached lattice)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench_configs as bc
import depth_estimation_b200 as pkg
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
Nn = pkg._native
for (h, w) in [(1024, 1024)]:
    feats = bc._features(h, w, 1, dev)
    lat = pkg.PermutohedralLattice(feats)
    n = h * w
    for C in (1, 3, 5, 16, 32, 64):
        src = torch.rand((n, C), device=dev)
        out = torch.empty_like(src)
        res = {}
        for name, flat in (("tiled", False), ("flat", True)):
            for _ in range(2):
                lat.filter(src, out=out, flat=flat)
            res[name] = bc._time(lambda: lat.filter(src, out=out, flat=flat), 7, flush, dev)
        Nn.profile(True)
        lat.filter(src, out=out)
        torch.cuda.synchronize()
        Nn.profile(False)
        rep = {k: round(v[1], 4) for k, v in Nn.profile_report().items()}
        print(json.dumps({"C": C, "ms": res, "tiled_kernels": rep}), file=sys.stderr)

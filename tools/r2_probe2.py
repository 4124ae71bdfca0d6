"""Round-2 probe: frame and filter timings on realistic lattices, per-kernel (scratch driver)."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import depth_estimation_b200 as pkg
import bench_util as U
import bench_extra as X
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
which = sys.argv[1:] or ["crf"]
out = []
if "crf" in which:
    for (h, w, L, v) in [(288, 384, 64, "sharp"), (500, 740, 64, "default"), (500, 740, 64, "sharp"),
                         (500, 740, 64, "sharp_textured"), (1080, 1920, 128, "sharp_textured")]:
        out.append(X.crf_entry(pkg, dev, 0, flush, h, w, L, v, 5, None))
if "filter" in which:
    for v in ("sharp", "sharp_textured"):
        out += X.filter_entries(pkg, dev, 0, flush, 1024, 1024, [64, 256], v, None)
for e in out:
    print(json.dumps({k: e[k] for k in ("workload", "value", "vertices", "M_over_n", "kernels_ms") if k in e}
                     | {"frac": e["roofline"]["frac"], "ms": e.get("ms_per_frame", e.get("ms_apply"))}), flush=True)

"""A/B of the small-lattice blur forms (PL_BLUR_FORM, read per call: "reg" = grid-barrier form, "flow" = default):
bitwise comparison of a filter apply and of a 5-iteration mean-field frame, then device time per frame.

    python tools/blur_exp.py [H W L variant]
"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import depth_estimation_b200 as pkg
import bench_util as U
synthetic = importlib.import_module("depth_estimation_b200.synthetic")

h, w, L, variant = 500, 740, 64, "default"
if len(sys.argv) > 4:
    h, w, L, variant = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
modes = os.environ.get("BLUR_MODES", "reg,flow").split(",")
dev = torch.device("cuda", 0)
feats = torch.from_numpy(U.features(h, w, variant)).to(dev)
g = torch.Generator(device=dev).manual_seed(1)
e0 = torch.randn((h * w, L), generator=g, device=dev).abs_().mul_(4)
src = torch.rand((h * w, L), generator=g, device=dev)
sparse = torch.zeros_like(src)
sparse[::5003] = torch.rand((len(range(0, h * w, 5003)), L), generator=g, device=dev) + 0.5
mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).to(dev)
lat = pkg.PermutohedralLattice(feats)
print("M", lat.num_vertices, "n", h * w, "L", L, flush=True)


def frame():
    return lat.mean_field(e0, mu, 5, 1.0)


res = {}
for m in modes:
    os.environ["PL_BLUR_FORM"] = m
    f = lat.filter(src).clone()
    f2 = lat.filter(src).clone()          # second call: buffers in the state the first one left
    e = frame().clone()
    e2 = frame().clone()
    torch.cuda.synchronize()
    sp = lat.filter(sparse).clone()
    res[m] = (f, e, sp)
    print("mode", m, "run-to-run max|d| filter %.3g mean-field %.3g" % ((f - f2).abs().max().item(), (e - e2).abs().max().item()),
          "finite", bool(torch.isfinite(e).all()), "| sparse source repeatable", bool(torch.equal(sp, lat.filter(sparse))), flush=True)
    if m != modes[0]:
        f0, e0_, sp0 = res[modes[0]]
        print("  sparse source (deterministic splat) bit-equal to mode", modes[0], ":", bool(torch.equal(sp, sp0)),
              "max|d| %.3g" % (sp - sp0).abs().max().item(), "nonzero", int((sp != 0).sum().item()))
        print("  vs mode", modes[0], ": filter bit-equal", bool(torch.equal(f, f0)),
              "max|d| %.3g" % (f - f0).abs().max().item(),
              "| mean-field bit-equal", bool(torch.equal(e, e0_)), "max|d| %.3g" % (e - e0_).abs().max().item(),
              flush=True)

for rep in range(2):
    for m in modes:
        os.environ["PL_BLUR_FORM"] = m
        for _ in range(5):
            frame()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(50):
            frame()
        b.record()
        torch.cuda.synchronize()
        print("rep", rep, "mode", m, "mean-field frame (lattice cached) %.1f us" % (a.elapsed_time(b) * 20), flush=True)
lat.close()

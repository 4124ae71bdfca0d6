"""Whole frames (lattice build + 5 iterations, a new lattice per frame as in bench.py) under the blur / launch
knobs, back to back and with a synchronise per frame:  python tools/frame_exp.py"""
import importlib, os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1 and sys.argv[1] == "child":
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
    cached = pkg.PermutohedralLattice(feats)

    def frame():
        lat = pkg.PermutohedralLattice(feats)
        out = lat.mean_field(e0, mu, 5, 1.0)
        lat.close()
        return out

    def timed(fn, reps):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) * 1000 / reps

    res = {}
    for rep in range(2):
        res["build+loop back-to-back"] = timed(frame, 300)
        res["loop only (cached lattice)"] = timed(lambda: cached.mean_field(e0, mu, 5, 1.0), 300)
        res["build only"] = timed(lambda: pkg.PermutohedralLattice(feats).close(), 300)
    print(json.dumps(res))
else:
    for cfg in ["", "PL_NO_PDL=1", "PL_BLUR_FORM=reg", "PL_NO_PDL=1 PL_BLUR_FORM=reg"] + sys.argv[1:]:
        env = dict(os.environ)
        for kv in cfg.split():
            k, v = kv.split("=")
            env[k] = v
        out = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True)
        try:
            r = json.loads(out.stdout.strip().splitlines()[-1])
            print("%-32s" % (cfg or "default"), "  ".join("%s %.1f us" % kv for kv in r.items()), flush=True)
        except Exception:
            print(cfg, "FAILED", out.stdout[-300:], out.stderr[-300:])

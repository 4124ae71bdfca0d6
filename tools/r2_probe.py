"""Round-2 probe: how the kernels behave on realistic lattices (module default s = 0.1 vs the
reference's sharp sigma_p = .01, sigma_c = .125, gaussian_matrix.py:498-505).  Scratch driver."""
import importlib, json, sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import depth_estimation_b200 as pkg
syn = importlib.import_module("depth_estimation_b200.synthetic")
Nn = pkg._native
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, reps=5):
    ms = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    ms.sort()
    return ms[len(ms) // 2]

def prof(fn):
    torch.cuda.synchronize()
    Nn.profile(True); fn(); torch.cuda.synchronize(); Nn.profile(False)
    return {k: (c, round(t, 4)) for k, (c, t) in Nn.profile_report().items()}

SCALES = {"s0.1": (0.1, 0.1), "sharp": (0.01, 0.125)}
which = sys.argv[1:] or ["crf", "sweep", "big"]
imgs = {}
def feats(h, w, name):
    if (h, w) not in imgs:
        imgs[(h, w)] = syn.image(h, w, 0)
    sij, srgb = SCALES[name]
    return torch.from_numpy(syn.ijrgb_features(imgs[(h, w)], sij, srgb)).to(dev)

def tile_stats(lat):
    st = lat.export()
    vid = st["vid"].long()
    n = vid.shape[0]
    T = n // 128
    v = vid[:T * 128].reshape(T, 768).sort(dim=1).values
    U = (v[:, 1:] != v[:, :-1]).sum(1) + 1
    return {"U_mean": float(U.float().mean()), "U_max": int(U.max()), "U_p99": float(torch.quantile(U.float(), 0.99)),
            "frac_gt120": float((U > 120).float().mean()), "frac_gt160": float((U > 160).float().mean())}

if "crf" in which:
    for (h, w, L) in [(288, 384, 16), (288, 384, 64), (500, 740, 64)]:
        n = h * w
        e0 = torch.from_numpy(syn.unary(h, w, L, 0)).to(dev)
        mu = torch.from_numpy(syn.charb_matrix(L)).to(dev)
        for name in SCALES:
            f = feats(h, w, name)
            box = {}
            def build():
                if "lat" in box: box["lat"].close()
                box["lat"] = pkg.PermutohedralLattice(f)
            for _ in range(3): build()
            tb = timeit(build)
            lat = box["lat"]
            for _ in range(2): lat.mean_field(e0, mu, 5)
            tm = timeit(lambda: lat.mean_field(e0, mu, 5))
            pr = prof(lambda: (build(), box["lat"].mean_field(e0, mu, 5)))
            lat = box["lat"]
            print(json.dumps({"w": "crf", "h": h, "wd": w, "L": L, "scale": name, "M": lat.num_vertices,
                              "M_over_n": lat.num_vertices / n, "ms_build": tb, "ms_mf5": tm, "tiles": tile_stats(lat), "prof": pr}), flush=True)
            lat.close()

if "sweep" in which:
    for (h, w) in [(1024, 1024), (4096, 2048)]:
        n = h * w
        for name in SCALES:
            f = feats(h, w, name)
            box = {}
            def build():
                if "lat" in box: box["lat"].close()
                box["lat"] = pkg.PermutohedralLattice(f)
            for _ in range(3): build()
            tb = timeit(build)
            lat = box["lat"]
            M = lat.num_vertices
            ts = tile_stats(lat)
            pb = prof(build); lat = box["lat"]
            for C in [4, 64, 256]:
                src = torch.rand((n, C), device=dev); out = torch.empty_like(src)
                for _ in range(2): lat.filter(src, out=out)
                ta = timeit(lambda: lat.filter(src, out=out))
                pr = prof(lambda: lat.filter(src, out=out))
                nbytes = 8 * n * C + 96 * n + 56 * M * C + 48 * M
                print(json.dumps({"w": "sweep", "h": h, "wd": w, "C": C, "scale": name, "M": M, "M_over_n": M / n,
                                  "ms_build": tb, "ms_apply": ta, "frac_survey_bytes": nbytes / (ta * 1e-3) / 1e9 / 6557.8,
                                  "tiles": ts, "prof": pr, "prof_build": pb if C == 4 else None}), flush=True)
                del src, out
            lat.close()

if "big" in which:
    for (h, w, L) in [(1450, 2000, 256), (1080, 1920, 128)]:
        n = h * w
        g = torch.Generator(device=dev).manual_seed(1)
        e0 = torch.randn((n, L), generator=g, device=dev).abs_() * 4
        mu = torch.from_numpy(syn.charb_matrix(L)).to(dev)
        for name in SCALES:
            f = feats(h, w, name)
            lat = pkg.PermutohedralLattice(f)
            for _ in range(2): lat.mean_field(e0, mu, 5)
            tm = timeit(lambda: lat.mean_field(e0, mu, 5), 3)
            pr = prof(lambda: lat.mean_field(e0, mu, 5))
            M = lat.num_vertices
            print(json.dumps({"w": "big", "h": h, "wd": w, "L": L, "scale": name, "M": M, "M_over_n": M / n,
                              "ms_mf5": tm, "tiles": tile_stats(lat), "prof": pr}), flush=True)
            lat.close()

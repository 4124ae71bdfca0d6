"""Hammer compat_tc_kernel (L = 128 / 256 frames, and the raw transform with odd row counts) on one GPU."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, ctypes as C
import depth_estimation_b200 as pkg, bench_util as U
syn = importlib.import_module("depth_estimation_b200.synthetic")
N = pkg._native
dev = torch.device("cuda", 0)
secs = float(sys.argv[1]) if len(sys.argv) > 1 else 40
t_end = time.time() + secs
cases = []
for (h, w, L) in [(1080, 1920, 128), (1450, 2000, 256)]:
    f = torch.from_numpy(U.features(h, w, "default")).to(dev)
    g = torch.Generator(device=dev).manual_seed(1)
    e0 = torch.randn((h * w, L), generator=g, device=dev).abs_().mul_(4)
    mu = torch.from_numpy(syn.charb_matrix(L, 0.05)).to(dev)
    cases.append((pkg.PermutohedralLattice(f), e0, mu))
frames = 0
while time.time() < t_end:
    for lat, e0, mu in cases:
        lat.mean_field(e0, mu, 5, 1.0)
        frames += 1
    # raw transform on ragged row counts / label shards
    for n, L, l0, cnt in [(100003, 256, 64, 64), (77777, 128, 0, 128), (1449888, 256, 0, 256)]:
        e = cases[1][1][:n, :L].contiguous() if L == 256 else cases[0][1][:n, :L].contiguous()
        mu_t = (cases[1][2] if L == 256 else cases[0][2]).t().contiguous()
        p = torch.empty((n, cnt), device=dev)
        N.check(N.lib().pl_softmax_compat_blocked(C.c_void_p(e.data_ptr()), 1, C.c_void_p(mu_t.data_ptr()),
                                                  C.c_void_p(p.data_ptr()), n, L, l0, cnt, C.c_float(1.0), N.stream_ptr(dev)))
    torch.cuda.synchronize()
print("stress_compat ok:", frames, "frames,", N.lib().pl_debug_tc_timeouts(), "timeouts")

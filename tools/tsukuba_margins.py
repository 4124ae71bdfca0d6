"""Margins of tests/test_fullsize_gpu.py::test_tsukuba_fixture_pair_to_depth_vs_reference_pipeline over
repeated runs of the GPU side (the CPU side is computed once)."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import depth_estimation_b200 as native
from oracle import depth_oracle as DO
from oracle.lattice_oracle import Oracle, Reference
synthetic = importlib.import_module("depth_estimation_b200.synthetic")
U = importlib.import_module("depth_estimation_b200.crf.utils")
cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
G = os.path.join(ROOT, "tests", "golden")
left, right = U.read_image(os.path.join(G, "imL.png")), U.read_image(os.path.join(G, "imR.png"))
h, w, _ = left.shape
L = w // 6
vol = DO.disparity_badness(left, right)
e0 = vol.reshape(h * w, L).astype(np.float32)
ref = synthetic.ijrgb_features(left.astype(np.float32), 0.1, 0.1)
mu = synthetic.charb_matrix(L, 0.05)
backend = Reference() if Reference.available() else Oracle(faithful=True)
e = e0.copy()
for _ in range(5):
    x = -e.astype(np.float64); q = np.exp(x - x.max(1, keepdims=True)); q /= q.sum(1, keepdims=True)
    p = (q @ mu.T.astype(np.float64)).astype(np.float32)
    f = backend.filter(np.ascontiguousarray(p), ref)
    e = e0 + (f - p)
want_e = e
x = -want_e.astype(np.float64); q = np.exp(x - x.max(1, keepdims=True))
want_depth = (q / q.sum(1, keepdims=True) * np.arange(L)[None]).sum(1).reshape(h, w)
dl, dr = torch.from_numpy(left).float().cuda(), torch.from_numpy(right).float().cuda()
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 40):
    depth, label = cm.stereo_pair_depth(dl, dr, niters=5, gamma=0.05, s_ij=0.1, s_rgb=0.1, return_label=True)
    depth, label = depth.cpu().numpy(), label.cpu().numpy()
    agree = (label.reshape(-1) == want_e.argmin(1)).mean()
    dmax, dmean = np.abs(depth - want_depth).max(), np.abs(depth - want_depth).mean()
    lat = native.PermutohedralLattice(torch.from_numpy(ref).cuda())
    got_e = lat.mean_field(torch.from_numpy(e0).cuda(), torch.from_numpy(mu).cuda(), 5, 1.0).cpu().numpy()
    lat.close()
    emax = np.abs(got_e - want_e).max() / np.abs(want_e).max()
    flag = "" if (agree > 0.999 and dmax < 2e-2 and dmean < 1e-3) else "   <-- would fail"
    print("rep %2d  label agreement %.5f  depth max %.3g mean %.3g  E rel max %.3g%s" % (rep, agree, dmax, dmean, emax, flag), flush=True)

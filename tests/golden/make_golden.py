"""Generate the golden vectors in tests/golden/ by running the REAL reference Python modules.

TEST INFRASTRUCTURE.  Runs only where /root/reference exists (this container); the vectors it
writes (small .npz files, committed) travel to the GPU box.

    python tests/golden/make_golden.py

The reference's crf/gaussian_matrix.py cannot be imported as it stands:
  * line 8 imports the un-vendored `guided_filter_pytorch`      -> stub module (classes unused here)
  * line 15 JIT-builds ~/depth-estimation/crf/lattice/lite/lattice.cpp, which no longer compiles
    against torch 2.11 (tensorFromBlob)                          -> torch.utils.cpp_extension.load is
    patched to return a module whose `filter` is oracle/_ref/libref_lattice.so, i.e. the
    unmodified reference header permutohedral.h compiled in place (oracle/build_ref.py).
Everything else -- LatticeFilter / BatchedLatticeFilter forward and backward, BatchedAdjacency,
LatticeGaussian, RbfLaplacian(C), charb, potts, compatibility_matrix, mean_field_infer,
CRFasRNN.forward, ijrgbGuide, ijGuide -- is the reference's own code, executed on CPU.
`crf/mb_stereo_crf.py` needs the un-vendored `oil` package at import; `oil`, the data loader and
matplotlib are stubbed so that its nn.Module forwards (logits2average_depth, CRFdepthRefiner) run.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def import_reference():
    from oracle.lattice_oracle import Reference
    ref = Reference()

    def ref_filter(src, refm):
        out = ref.filter(src.detach().contiguous().numpy(), refm.detach().contiguous().numpy())
        return torch.from_numpy(out)

    class _Dummy(torch.nn.Module):
        def __init__(self, *a, **k):
            super().__init__()

    gfp = _stub("guided_filter_pytorch")
    gfp.guided_filter = _stub("guided_filter_pytorch.guided_filter", GuidedFilter=_Dummy,
                              BoxFilter=_Dummy)
    import torch.utils.cpp_extension as ce
    ce.load = lambda *a, **k: types.SimpleNamespace(filter=ref_filter)

    # stubs for crf/mb_stereo_crf.py's imports (oil, data loader, plotting)
    class _Trainer:
        def __init__(self, *a, **k):
            pass
    oil = _stub("oil")
    oil.utils = _stub("oil.utils")
    _stub("oil.utils.utils", Eval=None, cosLr=None, LoaderTo=None, Named=type)
    _stub("oil.model_trainers", Trainer=_Trainer, Classifier=_Trainer, Regressor=_Trainer)
    _stub("oil.utils.mytqdm", tqdm=lambda x, **k: x)
    _stub("oil.architectures")
    _stub("oil.architectures.parts")
    class _CoordConv(torch.nn.Module):
        """Stand-in for oil's un-vendored ``conv2d(cin, cout, coords=True)`` (mb_stereo_crf.py:5-7,
        :84-90): 3x3 same-padded convolution over the input plus normalised (i, j) channels.
        The product's ``_CoordConv2d`` is the same re-creation; the golden pins everything around it."""
        def __init__(self, cin, cout, coords=False, **k):
            super().__init__()
            self.conv = torch.nn.Conv2d(cin + 2, cout, kernel_size=3, padding=1)

        def forward(self, x):
            b, _, h, w = x.shape
            ii = torch.linspace(-1, 1, h)[None, None, :, None].expand(b, 1, h, w)
            jj = torch.linspace(-1, 1, w)[None, None, None, :].expand(b, 1, h, w)
            return self.conv(torch.cat([x, ii, jj], dim=1))
    _stub("oil.architectures.parts.blocks", conv2d=_CoordConv, ConvBNrelu=None, ResBlock=None)
    _stub("crf.dataloader", MBStereo14Unary=None, StereoUpsampling05=None)
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib  # noqa: F401
        except Exception:
            mpl = _stub("matplotlib")
            mpl.pyplot = _stub("matplotlib.pyplot")

    sys.path.insert(0, REF)
    import crf.gaussian_matrix as rgm
    import crf.crf_module as rcm
    # the reference fans the batch out over mp.Pool (gaussian_matrix.py:370-377); same filter,
    # run in-process here
    rgm.batched_filter = lambda srcs, refs, nt: torch.stack(
        [rgm.latticefilter(s.contiguous(), r.contiguous()) for s, r in zip(srcs, refs)])
    try:
        import crf.mb_stereo_crf as rmb
    except Exception as exc:  # pragma: no cover - informational
        print("crf.mb_stereo_crf not importable:", exc)
        rmb = None
    return ref, rgm, rcm, rmb


def stereo_modules(synthetic, rgm, rcm, rmb):
    """mb_stereo_crf.py:62-102,138-163: the three nn.Module forwards with the lattice pairwise
    (``CRF.W = BatchedAdjacency``, the drop-in crf_module.py:5 imports), on CPU.  The reference
    hard-codes ``.cuda()`` for the labels (:150; crf_module.py:73): Tensor.cuda is a no-op here."""
    g = torch.Generator().manual_seed(21)
    torch.manual_seed(21)
    torch.Tensor.cuda = lambda self, *a, **k: self
    rec = {}
    lg = torch.randn((2, 5, 6, 7), generator=g)
    rec.update(logits=lg.numpy(), depth=rmb.logits2average_depth(lg).numpy(),
               depth_labels=rmb.logits2average_depth(lg, torch.linspace(0, 2, 5)[None, :, None, None]).numpy())

    def state(m, prefix):
        return {prefix + k.replace(".", "__"): v.detach().numpy() for k, v in m.state_dict().items()}

    # CRFdepthRefiner: 16-channel guide (rgb + 13 projected features) -> d = 16 lattice, L = 16 labels
    h, w, L, d_in = 14, 18, 16, 8
    img = torch.from_numpy(synthetic.image(h, w, seed=31)).permute(2, 0, 1)[None].contiguous()
    feats = torch.randn((1, d_in, h, w), generator=g) * 0.5
    logits = (-torch.from_numpy(synthetic.unary(h, w, L, 31)).reshape(h, w, L)).permute(2, 0, 1)[None].contiguous()
    with torch.no_grad():
        m = rmb.CRFdepthRefiner(d_in=d_in, d_guide=16, niters=2, gamma=.05)
        m.CRF.W = rgm.BatchedAdjacency(num_threads=1)
        m.projection.weight.mul_(4.0)          # spread the projected features over a few lattice cells
        out = m((logits, img * 8.0, feats))
        rec.update(state(m, "refiner__"), refiner_logits=logits.numpy(), refiner_img=(img * 8.0).numpy(),
                   refiner_feats=feats.numpy(), refiner_out=out.numpy())
        # CRFwUncertainty: same plus the confidence map exp(-s)
        m = rmb.CRFwUncertainty(d_in=d_in, d_guide=16, niters=2, gamma=.05)
        m.CRF.W = rgm.BatchedAdjacency(num_threads=1)
        m.projection.weight.mul_(4.0)
        out, conf = m((logits, img * 8.0, feats))
        rec.update(state(m, "uncert__"), uncert_out=out.numpy(), uncert_conf=conf.numpy())
        # CRFdepthUpsampler: 18 linspace labels from the low-resolution disparity (:138-163)
        h2, w2 = 24, 32
        img2 = torch.from_numpy(synthetic.image(h2, w2, seed=33)).permute(2, 0, 1)[None].contiguous()
        disp = torch.from_numpy((np.add.outer(np.arange(h2 // 4), 1.5 * np.arange(w2 // 4)) % 7 + 0.5)
                                .astype(np.float32))[None, None]
        disp[0, 0, 0, :2] = 0.0                # unknown disparities: confidence 0 (:152)
        m = rmb.CRFdepthUpsampler(niters=2, gamma=.05)
        m.CRF.W = rgm.BatchedAdjacency(num_threads=1)
        out = m((disp, img2 * 10.0, None))
        rec.update(state(m, "upsampler__"), upsampler_disp=disp.numpy(), upsampler_img=(img2 * 10.0).numpy(),
                   upsampler_out=out.numpy())
        # a second frame with a different max_depth through the SAME module (label-cache regression)
        out2 = m((disp * 1.7, img2 * 10.0, None))
        rec.update(upsampler_out2=out2.numpy())
    np.savez_compressed(os.path.join(HERE, "stereo_wrappers.npz"), **rec)


def main():
    synthetic = __import__("importlib").import_module("depth_estimation_b200.synthetic")
    ref, rgm, rcm, rmb = import_reference()
    torch.manual_seed(0)
    out = {}
    if "--only-stereo" in sys.argv:
        stereo_modules(synthetic, rgm, rcm, rmb)
        print("stereo_wrappers.npz written")
        return

    # ---- 1. lattice integer artefacts + filter on an (r,g,b,i,j)-ordered guide -------------
    h, w = 24, 32
    img = synthetic.image(h, w, seed=5)
    feats = synthetic.ijrgb_features(img, 0.05, 0.125, rgb_first=True)   # gaussian_matrix.py:498-505
    st = ref.state(feats)
    src = np.random.default_rng(1).random((h * w, 7)).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "lattice_state_24x32.npz"), ref=feats, src=src,
                        greedy=st.greedy, rank=st.rank, bary=st.bary, vid=st.vid, keys=st.keys,
                        nbr=st.nbr, M=np.int64(st.M), filtered=ref.filter(src, feats))

    # ---- 2. LatticeFilter forward / backward (gaussian_matrix.py:423-468) ------------------
    g = torch.Generator().manual_seed(2)
    n, L, d = 300, 3, 3
    s = torch.rand((n, L), generator=g).requires_grad_()
    r = (torch.rand((n, d), generator=g) * 3).requires_grad_()
    gout = torch.randn((n, L), generator=g)
    y = rgm.LatticeFilter.apply(s, r)
    y.backward(gout)
    np.savez_compressed(os.path.join(HERE, "lattice_filter_autograd.npz"), src=s.detach().numpy(),
                        ref=r.detach().numpy(), gout=gout.numpy(), out=y.detach().numpy(),
                        grad_src=s.grad.numpy(), grad_ref=r.grad.numpy())

    # ---- 3. operators on one lattice: LatticeGaussian, RbfLaplacian(C) (:292-338) ----------
    # guide scales 0.2: every point has neighbours (degree K1 - 1 >= 2), so the normalised
    # Laplacians are well conditioned (with 0.1 some degrees are ~0 or negative: 'sym' is NaN)
    refm = torch.from_numpy(synthetic.ijrgb_features(synthetic.image(20, 28, seed=7), 0.2, 0.2))
    U = torch.rand((refm.shape[0], 2), generator=g)
    ops = {"gaussian": rgm.LatticeGaussian(refm) @ U,
           "lap_norm": rgm.RbfLaplacian(refm, normalize=True) @ U,
           "lap_unnorm": rgm.RbfLaplacian(refm, normalize=False) @ U,
           "lapc_sym": rgm.RbfLaplacianC(refm, 'sym') @ U,
           "lapc_right": rgm.RbfLaplacianC(refm, 'right') @ U,
           "lapc_none": rgm.RbfLaplacianC(refm, None) @ U}
    np.savez_compressed(os.path.join(HERE, "lattice_operators.npz"), ref=refm.numpy(), U=U.numpy(),
                        **{k: v.detach().numpy() for k, v in ops.items()})

    # ---- 4. CRFasRNN forward + backward with W = BatchedAdjacency (crf_module.py:81-104) ----
    h, w, L, niters = 20, 28, 8, 3
    img = synthetic.image(h, w, seed=9)
    rgb = torch.from_numpy(img).permute(2, 0, 1)[None].contiguous()
    logits = (-torch.from_numpy(synthetic.unary(h, w, L, 9)).reshape(h, w, L)).permute(2, 0, 1)[None]
    logits = logits.contiguous().requires_grad_()
    guide_mod = rcm.ijrgbGuide(0.1, 0.1)
    crf = rcm.CRFasRNN(rcm.charb(0.05), niters=niters)
    crf.W = rgm.BatchedAdjacency(num_threads=1)     # the lattice pairwise (imported at crf_module.py:5)
    labels = torch.arange(L).float()                 # charb's default is .cuda() (crf_module.py:73)
    conf = torch.rand((1, 1, h, w), generator=g) * 0.5 + 0.75
    guide = guide_mod(rgb)
    out_logits = crf(guide, logits, confidence=conf, labels=labels)
    target = torch.from_numpy(synthetic.disparity(h, w, L, 9)).float()[None, None]
    probs = torch.softmax(out_logits, dim=1)
    depth = (probs * labels[None, :, None, None]).sum(1)[:, None]
    loss = (depth - target).abs().mean()
    loss.backward()
    np.savez_compressed(os.path.join(HERE, "crf_as_rnn_20x28.npz"), rgb=rgb.numpy(),
                        logits=logits.detach().numpy(), confidence=conf.numpy(),
                        guide=guide.detach().numpy(), out_logits=out_logits.detach().numpy(),
                        depth=depth.detach().numpy(), loss=np.float32(loss.item()),
                        target=target.numpy(), grad_logits=logits.grad.numpy(),
                        grad_gamma=crf.Mu.gamma.grad.numpy(), grad_s=crf.Mu.s.grad.numpy(),
                        grad_s_ij=guide_mod.s_ij.grad.numpy(), grad_s_rgb=guide_mod.s_rgb.grad.numpy(),
                        niters=np.int64(niters), gamma=np.float32(0.05))

    # ---- 5. matrix form: mean_field_infer, compatibility_matrix, potts, charb energies -----
    n, L = refm.shape[0], 6
    E0 = torch.rand((n, L), generator=g) * 4
    lab = torch.linspace(0, 3, L)
    Mu = rcm.compatibility_matrix(rcm.charbonneir, lab)
    Q = rcm.mean_field_infer(E0, rgm.LatticeGaussian(refm), Mu, niters=4)
    pt = rcm.potts(5)
    x = torch.rand((1, 5, 6, 7), generator=g)
    ch = rcm.charb(0.05)
    disp = torch.rand((1, 1, 6, 7), generator=g) * 3
    en = ch.get_energies_from_scalar(disp, lab[None, :, None, None])
    np.savez_compressed(os.path.join(HERE, "crf_matrix_form.npz"), ref=refm.numpy(), E0=E0.numpy(),
                        labels=lab.numpy(), Mu=Mu.numpy(), Q=Q.detach().numpy(), potts_x=x.numpy(),
                        potts_y=pt(x).detach().numpy(), disp=disp.numpy(),
                        energies=en.detach().numpy(),
                        Mu2=rcm.compatibility_matrix(rcm.charbonneir2, lab).numpy(),
                        ij_guide=rcm.ijGuide(0.2)(x).detach().numpy())

    # ---- 6. stereo wrappers and Module forwards (mb_stereo_crf.py:62-102,138-163) -----------
    if rmb is not None:
        stereo_modules(synthetic, rgm, rcm, rmb)
    # ---- 7. cost volume and fixture readers: crf/depth.py:34-53, crf/utils.py:46-109 --------
    # crf/depth.py imports cv2.ximgproc (opencv-contrib, absent) for an unrelated guided-filter
    # operator and crf/utils.py imports skimage for `grayscale`: both stubbed, neither is called.
    import cv2
    if not hasattr(cv2, "ximgproc"):
        cv2.ximgproc = _stub("cv2.ximgproc", guidedFilter=None)
    if "skimage" not in sys.modules:
        sk = _stub("skimage")
        sk.color = _stub("skimage.color", rgb2gray=None)
    import crf.depth as rdepth
    import crf.utils as rutils
    left = synthetic.image(26, 44, seed=3).astype(np.float64)
    disp = synthetic.disparity(26, 44, 44 // 6, 3)
    right = np.zeros_like(left)
    for y in range(26):
        for x in range(44):
            xs = x + int(disp[y, x])
            if xs < 44:
                right[y, x] = left[y, xs]
    np.savez_compressed(os.path.join(HERE, "cost_volume_26x44.npz"), left=left, right=right,
                        ad9=rdepth.disparity_badness(left, right),
                        sd5=rdepth.disparity_badness(left, right, 5, rdepth.SD),
                        np3=rdepth.disparity_badness(left, right, 3, rdepth.nprod),
                        est=rdepth.disparity_estimate(left, right))
    import tempfile
    rng = np.random.default_rng(4)
    with tempfile.TemporaryDirectory() as tmp:
        gray = rng.integers(0, 256, (9, 13), dtype=np.uint8)
        pgm = b"P5\n# synthetic ground truth\n13 9\n255\n" + gray.tobytes()
        with open(os.path.join(tmp, "a.pgm"), "wb") as f:
            f.write(pgm)
        pf = rng.random((7, 11)).astype("<f4")
        pfm_le = b"Pf\n11 7\n-1.0\n" + pf.tobytes()
        pfm_be = b"Pf\n11 7\n1.0\n" + pf.astype(">f4").tobytes()
        for name, blob in (("le.pfm", pfm_le), ("be.pfm", pfm_be)):
            with open(os.path.join(tmp, name), "wb") as f:
                f.write(blob)
        from PIL import Image
        rgb8 = rng.integers(0, 256, (6, 8, 3), dtype=np.uint8)
        Image.fromarray(rgb8).save(os.path.join(tmp, "c.png"))
        with open(os.path.join(tmp, "c.png"), "rb") as f:
            png = f.read()
        np.savez_compressed(os.path.join(HERE, "fixture_readers.npz"),
                            pgm_bytes=np.frombuffer(pgm, np.uint8), pgm=rutils.read_pgm(os.path.join(tmp, "a.pgm")),
                            pfm_le_bytes=np.frombuffer(pfm_le, np.uint8),
                            pfm_le=rutils.read_pfm(os.path.join(tmp, "le.pfm")),
                            pfm_be_bytes=np.frombuffer(pfm_be, np.uint8),
                            pfm_be=rutils.read_pfm(os.path.join(tmp, "be.pfm")),
                            png_bytes=np.frombuffer(png, np.uint8),
                            png=rutils.read_image(os.path.join(tmp, "c.png")))
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()

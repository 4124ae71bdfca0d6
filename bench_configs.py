"""Measurements of the other BASELINE.json configurations (bench.py keeps the driver contract on
configs[1]; these are the parity-test shapes timed for profiles/ and DESIGN.md).

    python bench_configs.py --workload filter_sweep                        # configs[2], 1 GPU
    python bench_configs.py --workload crf_full_L256 [--steps K]           # configs[3], 1 GPU
    torchrun --nproc-per-node G ... bench_configs.py --workload crf_full_L256   # label-sharded, NCCL
    torchrun --nproc-per-node G ... bench_configs.py --workload crf_1080p_L128  # configs[4], image-parallel

One JSON line per measurement (rank 0).  Timing: CUDA events on the launching stream, inputs
resident in HBM and larger than L2 (or L2 flushed between repetitions), max over ranks.
"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

D = 5


def _peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def _features(h, w, seed, dev):
    """ijrgbGuide features [n, 5] (crf_module.py:111-114) of a smooth synthetic image, on the device."""
    import numpy as np
    import torch
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    # image() is O(n) numpy with 18 cosine terms: build a half-resolution image and upsample
    hh, ww = max(8, h // 4), max(8, w // 4)
    img = torch.from_numpy(synthetic.image(hh, ww, seed)).permute(2, 0, 1)[None].to(dev)
    img = torch.nn.functional.interpolate(img, size=(h, w), mode="bilinear", align_corners=False)
    g = torch.Generator(device=dev).manual_seed(seed)
    img = (img + 0.05 * torch.rand(img.shape, generator=g, device=dev)).clamp_(0, 1)[0]
    ii, jj = torch.meshgrid(torch.arange(h, device=dev), torch.arange(w, device=dev), indexing="ij")
    diag = float(np.sqrt(h * h + w * w))
    feats = torch.stack([ii / diag / 0.1, jj / diag / 0.1, img[0] / 0.1, img[1] / 0.1, img[2] / 0.1], -1)
    return feats.reshape(h * w, 5).float().contiguous()


def _time(fn, reps, flush, dev):
    import torch
    ms = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize(dev)
        ms.append(a.elapsed_time(b))
    ms.sort()
    return ms[len(ms) // 2]


def filter_sweep(args, dev):
    import torch
    import depth_estimation_b200 as pkg
    peak, how = _peak()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    shapes = [(1024, 1024), (2048, 1024), (2048, 2048), (4096, 2048)]
    chans = [1, 4, 16, 64, 256]
    if args.quick:
        shapes, chans = shapes[:2], [4, 64]
    for (h, w) in shapes:
        n = h * w
        feats = _features(h, w, 1, dev)
        box = {}

        def build():
            if "lat" in box:
                box["lat"].close()
            box["lat"] = pkg.PermutohedralLattice(feats)
        for _ in range(4):     # the stream-ordered pool grows during the first builds of a new size
            build()
        t_build = _time(build, 5, flush, dev)
        lat = box["lat"]
        M = lat.num_vertices
        for C in chans:
            src = torch.rand((n, C), device=dev)
            out = torch.empty_like(src)
            for _ in range(2):
                lat.filter(src, out=out)
            t_apply = _time(lambda: lat.filter(src, out=out), 5, flush, dev)
            # src + dst, per-entry lists, splat write + 2(d+1) blur + slice read
            nbytes = 8 * n * C + 15 * n * (D + 1) + 56 * M * C + 48 * M
            print(json.dumps({
                "metric": "lattice_filter_mpix_labels_per_s", "unit": "Mpix*labels/s",
                "value": n * C / (t_apply * 1e-3) / 1e6,
                "value_incl_build": n * C / ((t_apply + t_build) * 1e-3) / 1e6,
                "config": {"workload": f"filter_sweep_{h}x{w}_C{C}", "n": n, "channels": C, "d": D,
                           "vertices": M, "l2": "flushed between repetitions"},
                "ms_apply": t_apply, "ms_build": t_build,
                "roofline": {"bound": "hbm", "achieved": nbytes / (t_apply * 1e-3) / 1e9, "peak": peak,
                             "unit": "GB/s", "frac": nbytes / (t_apply * 1e-3) / 1e9 / peak,
                             "peak_source": how, "algorithmic_bytes": nbytes,
                             "scope": "whole filter apply (splat + blur + slice)"}}), flush=True)
            del src, out
        lat.close()


def crf_train_step(args, dev):
    """BASELINE config 2 as the reference trains it: CRFasRNN forward + backward through the
    autograd operators (LatticeFilter.backward, gaussian_matrix.py:441-468), gradients wrt the
    logits, gamma, s (compatibility) and s_ij, s_rgb (guide).  Also the cheaper step with the
    guide frozen (no reference gradient: backward = one filter per iteration)."""
    import time
    import torch
    import depth_estimation_b200 as pkg
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    h, w, L, niters = (125, 185, 64, 5) if args.quick else (500, 740, 64, 5)
    img = torch.from_numpy(synthetic.image(h, w, 0)).permute(2, 0, 1)[None].contiguous().to(dev)
    logits0 = (-torch.from_numpy(synthetic.unary(h, w, L, 0))).reshape(h, w, L).permute(2, 0, 1)[None]
    logits0 = logits0.contiguous().to(dev)
    target = torch.from_numpy(synthetic.disparity(h, w, L, 0)).float()[None, None].to(dev)
    labels = torch.arange(L, dtype=torch.float32, device=dev)
    for name, train_guide in (("guide_trainable", True), ("guide_frozen", False)):
        guide_mod = cm.ijrgbGuide(0.1, 0.1, trainable=train_guide).to(dev)
        crf = cm.CRFasRNN(cm.charb(0.05), niters=niters).to(dev)

        def step():
            lg = logits0.clone().requires_grad_()
            out = crf(guide_mod(img), lg)
            depth = (torch.softmax(out, dim=1) * labels[None, :, None, None]).sum(1)[:, None]
            loss = (depth - target).abs().mean()
            loss.backward()
            gm.clear_lattice_cache()
            return loss
        for _ in range(2):
            step()
        torch.cuda.synchronize(dev)
        ms = []
        for _ in range(max(3, args.steps)):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step()
            b.record()
            torch.cuda.synchronize(dev)
            ms.append(a.elapsed_time(b))
        ms.sort()
        med = ms[len(ms) // 2]
        peak_mem = torch.cuda.max_memory_allocated(dev) / 2 ** 30
        print(json.dumps({"metric": "stereo_crf_train_steps_per_s", "unit": "steps/s", "value": 1e3 / med,
                          "ms_per_step": med,
                          "config": {"workload": f"crf_train_{h}x{w}_L{L}_{niters}mf_{name}", "h": h, "w": w,
                                     "labels": L, "mf_iters": niters,
                                     "grads": "logits, gamma, s" + (", s_ij, s_rgb" if train_guide else "")},
                          "peak_mem_gib": peak_mem}), flush=True)


def ref_cuda_filter(args, dev):
    """The reference's own CUDA lattice (crf/lattice/cuda_lattice, rebuilt in place for sm_100a by
    oracle/build_ref.py) beside ours on the shapes it supports: vd = 3 value channels, pd = 5,
    host buffers in and out (its MirroredArray copies are part of its call).  Speed only: it is
    Adams' normalised filter, not the lite path's arithmetic."""
    import ctypes as C
    import time
    import numpy as np
    import torch
    import depth_estimation_b200 as pkg
    so = os.path.join(ROOT, "oracle", "_ref", "libref_cuda_lattice.so")
    if not os.path.exists(so):
        print(json.dumps({"workload": "ref_cuda_filter", "unavailable": "oracle/_ref/libref_cuda_lattice.so not built"}))
        return
    lib = C.CDLL(so)
    lib.filter.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_bool]
    lib.filter.restype = None
    N = pkg._native
    for (h, w) in ([(1024, 1024)] if args.quick else [(500, 740), (1024, 1024), (2048, 1024)]):
        n = h * w
        feats = _features(h, w, 1, dev).cpu().numpy()
        src = np.random.default_rng(0).random((n, 3)).astype(np.float32)
        # reference CUDA: filters `im` in place (host pointers)
        t_ref = []
        for rep in range(1 + max(1, args.steps)):
            im = src.copy()
            t0 = time.perf_counter()
            lib.filter(im.ctypes.data, feats.ctypes.data, 5, 3, w, h, True)
            torch.cuda.synchronize(dev)
            t_ref.append(time.perf_counter() - t0)
        t_ref = sorted(t_ref[1:])[len(t_ref[1:]) // 2]
        # ours through the same kind of call: host buffers in, host buffer out (pl_filter_host)
        out = np.empty_like(src)
        t_our = []
        for rep in range(1 + max(1, args.steps)):
            t0 = time.perf_counter()
            N.check(N.lib().pl_filter_host(C.c_void_p(src.ctypes.data), C.c_void_p(feats.ctypes.data),
                                           C.c_void_p(out.ctypes.data), n, 3, 5, C.c_void_p(0)))
            t_our.append(time.perf_counter() - t0)
        t_our = sorted(t_our[1:])[len(t_our[1:]) // 2]
        # ours with device-resident buffers (what the CRF loop sees), lattice build included
        fd, sd = torch.from_numpy(feats).to(dev), torch.from_numpy(src).to(dev)
        od = torch.empty_like(sd)
        for _ in range(2):
            pkg.lattice.filter(sd, fd)
        t_dev = _time(lambda: pkg.lattice.filter(sd, fd), 5, None, dev) * 1e-3
        print(json.dumps({
            "metric": "lattice_filter_mpix_labels_per_s", "unit": "Mpix*labels/s",
            "config": {"workload": f"ref_cuda_filter_{h}x{w}_C3", "n": n, "channels": 3, "d": 5,
                       "call": "host buffers in/out, lattice built inside the call"},
            "reference_cuda": {"value": n * 3 / t_ref / 1e6, "ms": t_ref * 1e3,
                               "source": "crf/lattice/cuda_lattice/permutohedral.cu rebuilt for sm_100a"},
            "ours_host_call": {"value": n * 3 / t_our / 1e6, "ms": t_our * 1e3},
            "ours_device_resident": {"value": n * 3 / t_dev / 1e6, "ms": t_dev * 1e3}}), flush=True)


def crf_big(args, dev, rank, world, h, w, L, frames_total, name, sharded):
    import torch
    import torch.distributed as dist
    import depth_estimation_b200 as pkg
    ls = importlib.import_module("depth_estimation_b200.crf.label_shards")
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    n = h * w
    feats = _features(h, w, 7 if sharded else 7 + rank, dev)     # sharded: every rank the same guide
    mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).to(dev)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    if sharded and world > 1:
        l0, cnt = ls.label_shard(L, world, rank)
        e0 = (torch.randn((n, cnt), generator=g, device=dev).abs_() * 4)
    else:
        e0 = (torch.randn((n, L), generator=g, device=dev).abs_() * 4)
    frames = 1 if sharded else max(1, frames_total // world)
    state = {}

    peer = {}

    def step():
        for _ in range(frames):
            lat = pkg.PermutohedralLattice(feats)
            if sharded and world > 1 and args.peer:
                if "b" not in peer:
                    peer["b"] = ls.PeerMemoryBackend(lat, n, e0.shape[1])
                peer["b"].lat = lat
                out = peer["b"].run(e0, mu, 5, 1.0).clone()
            elif sharded and world > 1:
                out = ls.LabelShardedMeanField(ls.CudaBackend(lat)).run(e0, mu, 5, 1.0)
            else:
                out = lat.mean_field(e0, mu, 5, 1.0)
            state["M"] = lat.num_vertices
            state["out"] = out
            lat.close()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(args.warmup):
        step()
    barrier()
    total = 0.0
    for _ in range(args.steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        a.record()
        step()
        b.record()
        torch.cuda.synchronize(dev)
        total += a.elapsed_time(b)
    if world > 1:
        t = torch.tensor([total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total = float(t.item())
    ms = total / args.steps
    nframes = frames if sharded else frames * world
    if args.profile and rank == 0:
        Nn = pkg._native
        Nn.profile(True)
        step()
        torch.cuda.synchronize(dev)
        Nn.profile(False)
        for k, (c, t) in sorted(Nn.profile_report().items(), key=lambda kv: -kv[1][1]):
            print(f"  {k:45s} x{c:3d} {t:9.3f} ms", file=sys.stderr)
    if rank == 0:
        peak, how = _peak()
        M = state["M"]
        # per frame: lattice build + 5 x (slice 3nL + compat 2nL + splat nL floats) on this path
        per_frame = 5 * (6 * 4 * n * L + 15 * n * (D + 1) + 56 * M * L)
        print(json.dumps({
            "metric": "stereo_crf_frames_per_s_5mf", "unit": "frames/s", "value": nframes / (ms * 1e-3),
            "n_gpus": world, "ms_per_step": ms, "steps": args.steps, "warmup": args.warmup,
            "scaling": "strong" if sharded else "weak", "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "h": h, "w": w, "labels": L, "mf_iters": 5, "d": D,
                       "frames_per_step": nframes, "vertices": M,
                       "parallelism": ((f"label-sharded x{world} (peer-memory reads inside the compat kernel)"
                                        if args.peer else
                                        f"label-sharded x{world} (NCCL all-gather per iteration)")
                                       if sharded and world > 1 else f"image-parallel x{world}"),
                       "l2": "inputs larger than L2"},
            "roofline": {"bound": "hbm", "achieved": per_frame * nframes / world / (ms * 1e-3) / 1e9,
                         "peak": peak, "unit": "GB/s", "peak_source": how,
                         "frac": per_frame * nframes / world / (ms * 1e-3) / 1e9 / peak,
                         "scope": "whole frame, unfused large-L path: 6nL floats per iteration"}}),
              flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", required=True,
                    choices=["filter_sweep", "crf_full_L256", "crf_1080p_L128", "crf_tsukuba",
                             "ref_cuda_filter", "crf_train_step", "build_profile"])
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--profile", action="store_true", help="per-kernel CUDA-event times of one step")
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--peer", action="store_true",
                    help="label-sharded runs: read the peers' shards through CUDA IPC mappings")
    args = ap.parse_args()
    sys.stdout.flush()          # only JSON lines on stdout (NCCL's banner goes to stderr)
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.workload == "filter_sweep":
        if rank == 0:
            filter_sweep(args, dev)
    elif args.workload == "crf_train_step":
        if rank == 0:
            crf_train_step(args, dev)
    elif args.workload == "build_profile":
        if rank == 0:
            import depth_estimation_b200 as pkg
            Nn = pkg._native
            for (h, w) in [(2048, 2048), (4096, 2048)]:
                feats = _features(h, w, 1, dev)
                for _ in range(2):
                    pkg.PermutohedralLattice(feats).close()
                torch.cuda.synchronize(dev)
                Nn.profile(True)
                lat = pkg.PermutohedralLattice(feats)
                torch.cuda.synchronize(dev)
                Nn.profile(False)
                rep = {k: round(v[1], 3) for k, v in Nn.profile_report().items()}
                print(json.dumps({"workload": f"build_profile_{h}x{w}", "vertices": lat.num_vertices,
                                  "kernel_ms": rep}), flush=True)
                lat.close()
    elif args.workload == "ref_cuda_filter":
        if rank == 0:
            ref_cuda_filter(args, dev)
    elif args.workload == "crf_full_L256":
        crf_big(args, dev, rank, world, 1450, 2000, 256, 1, "middlebury_full_2000x1450_L256_5mf", True)
    elif args.workload == "crf_1080p_L128":
        crf_big(args, dev, rank, world, 1080, 1920, 128, args.frames or 8 * world,
                "batch_1080p_L128_5mf", False)
    else:
        crf_big(args, dev, rank, world, 288, 384, 16, args.frames or 8 * world,
                "tsukuba_384x288_L16_5mf", False)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""The other BASELINE.json configurations, measured inside the same bench.py invocation and reported
under `extra.configs` of its JSON line (bench.py's `value` stays on configs[1]):

  configs[0]  Tsukuba 384x288, 16 labels (and 64: max_disp = w // 6, crf/depth.py:40), 5 iterations
  configs[1]  740x500 x 64 on the sharp feature scales (the default scales are the headline)
  configs[2]  lattice filter sweep, {1, 8} MP x C in {4, 64, 256}: Mpix*labels/s
  configs[3]  2000x1450 x 256, one frame on one GPU; label-sharded over the ranks when N > 1
  configs[4]  1920x1080 x 128 frames, image-parallel over the ranks

Every entry carries its own NVML clock record, vertex count M and M/n, a whole-workload roofline
on SURVEY 8d's ALGORITHMIC bytes, the dominant kernel, and a CPU baseline extrapolated from a bounded
sample of the reference's own filter (labelled as extrapolated).  Each workload runs on the feature
variants of bench_util.VARIANTS (module default scales; the reference's sharp sigma_p = .01,
sigma_c = .125; the sharp scales on a heavily textured image)."""
import importlib
import time

import bench_util as U

D = U.D


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


def _profile(N, fn, dev):
    import torch
    torch.cuda.synchronize(dev)
    N.profile(True)
    fn()
    torch.cuda.synchronize(dev)
    N.profile(False)
    return {k: (c, t) for k, (c, t) in N.profile_report().items()}


def _top_kernel(rep, n, M, C, niters, peak):
    if not rep:
        return None
    name, (cnt, ms) = max(rep.items(), key=lambda kv: kv[1][1])
    total = sum(v[1] for v in rep.values())
    out = {"kernel": name, "launches": cnt, "ms": ms, "share": ms / total if total else None}
    nb = U.kernel_bytes(name, n, M, C, niters, cnt)
    if nb is not None and ms > 0:
        out["achieved_gbs"] = nb * cnt / (ms * 1e-3) / 1e9
        out["frac"] = out["achieved_gbs"] / peak
    return out


def _cpu_model(cpu_models, h, w, variant, feats_h):
    key = (h, w, variant)
    if key not in cpu_models:
        cpu_models[key] = U.CpuFilterModel(U.cpu_backend(), feats_h)
    return cpu_models[key]


def crf_entry(pkg, dev, local_rank, flush, h, w, L, variant, niters, cpu_models, reps=5):
    """One stereo-CRF frame (lattice build + `niters` mean-field iterations), inputs resident in HBM."""
    import torch
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    n = h * w
    feats_h = U.features(h, w, variant)
    feats = torch.from_numpy(feats_h).to(dev)
    if n * L <= (1 << 25):
        e0 = torch.from_numpy(synthetic.unary(h, w, L, 0)).to(dev)
    else:   # |N(0,1)| * 4 for >= 1 MP, drawn on the device
        g = torch.Generator(device=dev).manual_seed(1)
        e0 = torch.randn((n, L), generator=g, device=dev).abs_().mul_(4)
    mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).to(dev)
    state = {}

    def step():
        lat = pkg.PermutohedralLattice(feats)
        state["out"] = lat.mean_field(e0, mu, niters, 1.0)
        state["M"] = lat.num_vertices
        lat.close()

    for _ in range(3):
        step()
    sampler = U.ClockSampler(local_rank).start()
    ms = _time(step, reps, flush, dev)
    t_end = time.perf_counter() + 0.25      # keep the load under the sampler long enough for NVML
    while time.perf_counter() < t_end:
        step()
    torch.cuda.synchronize(dev)
    clocks = sampler.stop()
    M = state["M"]
    peak, how = U.peaks()
    rep = _profile(pkg._native, step, dev)
    nbytes = U.bytes_frame(n, M, L, niters)
    entry = {
        "workload": f"crf_{w}x{h}_L{L}_{niters}mf_{variant}", "metric": "stereo_crf_frames_per_s_5mf",
        "unit": "frames/s", "value": 1e3 / ms, "ms_per_frame": ms, "h": h, "w": w, "labels": L,
        "mf_iters": niters, "features": variant, **U.VARIANTS[variant], "vertices": M, "M_over_n": M / n,
        "mpix_labels_per_s_incl_build": n * L * niters / (ms * 1e-3) / 1e6,
        "l2": "flushed between timed frames (256 MiB write)", "clocks": clocks,
        "roofline": {"bound": "hbm", "scope": "whole frame: B_build + niters * B_iter (SURVEY 8d)",
                     "algorithmic_bytes": nbytes, "achieved": nbytes / (ms * 1e-3) / 1e9, "peak": peak,
                     "unit": "GB/s", "frac": nbytes / (ms * 1e-3) / 1e9 / peak, "peak_source": how,
                     "dominant_kernel": _top_kernel(rep, n, M, L, niters, peak)},
        "kernels_ms": {k: round(v[1], 4) for k, v in sorted(rep.items(), key=lambda kv: -kv[1][1])[:8]},
    }
    if cpu_models is not None:
        m = _cpu_model(cpu_models, h, w, variant, feats_h)
        secs = niters * (m.filter_seconds(n, L) + n * U.cpu_compat_seconds_per_row(L))
        entry["cpu_baseline"] = {"value": 1.0 / secs, "unit": "frames/s", "cores": 1, "kind": m.kind,
                                 "extrapolated": True,
                                 "sample": m.sample_text() + f"; + softmax / compat on a row sample; "
                                           f"x{niters} iterations = {secs:.1f} s per frame"}
    del e0, feats, state
    return entry


def filter_entries(pkg, dev, local_rank, flush, h, w, chans, variant, cpu_models, reps=5):
    """Lattice filter apply on a cached lattice (and including the build) for several channel counts."""
    import torch
    n = h * w
    feats_h = U.features(h, w, variant)
    feats = torch.from_numpy(feats_h).to(dev)
    box = {}

    def build():
        if "lat" in box:
            box["lat"].close()
        box["lat"] = pkg.PermutohedralLattice(feats)

    for _ in range(4):     # the stream-ordered pool grows during the first builds of a new size
        build()
    t_build = _time(build, reps, flush, dev)
    peak, how = U.peaks()
    rep_b = _profile(pkg._native, build, dev)
    lat = box["lat"]
    M = lat.num_vertices
    out = []
    for C in chans:
        src = torch.rand((n, C), device=dev)
        dst = torch.empty_like(src)
        for _ in range(2):
            lat.filter(src, out=dst)
        sampler = U.ClockSampler(local_rank).start()
        t_apply = _time(lambda: lat.filter(src, out=dst), reps, flush, dev)
        t_end = time.perf_counter() + 0.2
        while time.perf_counter() < t_end:
            lat.filter(src, out=dst)
        torch.cuda.synchronize(dev)
        clocks = sampler.stop()
        rep = _profile(pkg._native, lambda: lat.filter(src, out=dst), dev)
        nb, nb_b = U.bytes_apply(n, M, C), U.bytes_build(n, M)
        entry = {
            "workload": f"filter_{w}x{h}_C{C}_{variant}", "metric": "lattice_filter_mpix_labels_per_s",
            "unit": "Mpix*labels/s", "value": n * C / (t_apply * 1e-3) / 1e6,
            "value_incl_build": n * C / ((t_apply + t_build) * 1e-3) / 1e6,
            "ms_apply": t_apply, "ms_build": t_build, "h": h, "w": w, "channels": C, "features": variant,
            **U.VARIANTS[variant], "vertices": M, "M_over_n": M / n,
            "l2": "flushed between repetitions (256 MiB write)", "clocks": clocks,
            "roofline": {"bound": "hbm",
                         "scope": "filter apply, cached lattice: 8nC + 96n + 56MC + 48M (SURVEY 8d)",
                         "algorithmic_bytes": nb, "achieved": nb / (t_apply * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": nb / (t_apply * 1e-3) / 1e9 / peak, "peak_source": how,
                         "build_frac": nb_b / (t_build * 1e-3) / 1e9 / peak,
                         "dominant_kernel": _top_kernel(rep, n, M, C, 0, peak)},
            "kernels_ms": {k: round(v[1], 4) for k, v in sorted(rep.items(), key=lambda kv: -kv[1][1])[:6]},
            "build_kernels_ms": {k: round(v[1], 4)
                                 for k, v in sorted(rep_b.items(), key=lambda kv: -kv[1][1])[:9]},
        }
        if cpu_models is not None:
            m = _cpu_model(cpu_models, h, w, variant, feats_h)
            secs = m.filter_seconds(n, C)
            entry["cpu_baseline"] = {"value": n * C / secs / 1e6, "unit": "Mpix*labels/s", "cores": 1,
                                     "kind": m.kind, "extrapolated": True, "sample": m.sample_text()}
        out.append(entry)
        del src, dst
    lat.close()
    return out


def train_entries(pkg, dev, local_rank, steps=5):
    """BASELINE config 2 as the reference trains it: CRFasRNN forward + backward through the autograd
    operators (LatticeFilter.backward, gaussian_matrix.py:441-468) at 740x500 x 64, 5 iterations;
    gradients wrt the logits, gamma, s and -- `guide_trainable` -- the guide scales s_ij, s_rgb."""
    import torch
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    h, w, L, niters = 500, 740, 64, 5
    img = torch.from_numpy(synthetic.image(h, w, 0)).permute(2, 0, 1)[None].contiguous().to(dev)
    logits0 = (-torch.from_numpy(synthetic.unary(h, w, L, 0))).reshape(h, w, L).permute(2, 0, 1)[None]
    logits0 = logits0.contiguous().to(dev)
    target = torch.from_numpy(synthetic.disparity(h, w, L, 0)).float()[None, None].to(dev)
    labels = torch.arange(L, dtype=torch.float32, device=dev)
    out = []
    for name, train_guide in (("guide_trainable", True), ("guide_frozen", False)):
        guide_mod = cm.ijrgbGuide(0.1, 0.1, trainable=train_guide).to(dev)
        crf = cm.CRFasRNN(cm.charb(0.05), niters=niters).to(dev)

        def step():
            lg = logits0.clone().requires_grad_()
            res = crf(guide_mod(img), lg)
            depth = (torch.softmax(res, dim=1) * labels[None, :, None, None]).sum(1)[:, None]
            (depth - target).abs().mean().backward()
            gm.clear_lattice_cache()

        for _ in range(3):
            step()
        torch.cuda.synchronize(dev)
        sampler = U.ClockSampler(local_rank).start()
        ms = []
        for _ in range(steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step()
            b.record()
            torch.cuda.synchronize(dev)
            ms.append(a.elapsed_time(b))
        t_end = time.perf_counter() + 0.2
        while time.perf_counter() < t_end:
            step()
        torch.cuda.synchronize(dev)
        clocks = sampler.stop()
        ms.sort()
        med = ms[len(ms) // 2]
        out.append({"workload": f"crf_train_{w}x{h}_L{L}_{niters}mf_{name}", "metric": "stereo_crf_train_steps_per_s",
                    "unit": "steps/s", "value": 1e3 / med, "train_step_ms": med, "h": h, "w": w, "labels": L,
                    "mf_iters": niters, "features": "default", "vertices": None, "M_over_n": None,
                    "grads": "logits, gamma, s" + (", s_ij, s_rgb" if train_guide else ""), "clocks": clocks})
    return out


def single_gpu_extras(pkg, dev, local_rank, flush, with_cpu=True, quick=False):
    """All N = 1 entries.  ~40 s of GPU work, ~25 s of bounded CPU samples."""
    import torch
    cpu_models = {} if with_cpu else None
    out = []
    crf_shapes = [(288, 384, 16, ("default", "sharp")), (288, 384, 64, ("default", "sharp")),
                  (500, 740, 64, ("sharp", "sharp_textured")),
                  (1080, 1920, 128, ("default", "sharp_textured")),
                  (1450, 2000, 256, ("default", "sharp", "sharp_textured"))]
    sweep_shapes = [(1024, 1024), (4096, 2048)]
    chans = [4, 64, 256]
    if quick:
        crf_shapes, sweep_shapes, chans = crf_shapes[:3], sweep_shapes[:1], [4, 64]
    for (h, w, L, variants) in crf_shapes:
        for v in variants:
            out.append(crf_entry(pkg, dev, local_rank, flush, h, w, L, v, 5, cpu_models,
                                 reps=3 if h * w * L > (1 << 28) else 5))
            torch.cuda.empty_cache()
    for (h, w) in sweep_shapes:
        for v in ("default", "sharp", "sharp_textured"):
            out.extend(filter_entries(pkg, dev, local_rank, flush, h, w, chans, v, cpu_models))
            torch.cuda.empty_cache()
    out.extend(train_entries(pkg, dev, local_rank))
    torch.cuda.empty_cache()
    return out


def multi_gpu_extras(pkg, dev, rank, local_rank, world, steps=3, warmup=2):
    """N > 1: config 4 label-sharded over the ranks (strong scaling) in four exchange forms -- NCCL
    all-gather (the baseline north_star names), peer-memory reads, the pixel / label transpose over NCCL
    all-to-alls, and the transpose as one kernel over peer memory -- and config 5 image-parallel."""
    import torch
    import torch.distributed as dist
    ls = importlib.import_module("depth_estimation_b200.crf.label_shards")
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    out = []
    peak, how = U.peaks()

    def barrier():
        dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(step):
        for _ in range(warmup):
            step()
        barrier()
        total = 0.0
        for _ in range(steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            a.record()
            step()
            b.record()
            torch.cuda.synchronize(dev)
            total += a.elapsed_time(b)
        t = torch.tensor([total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    h, w, L, niters = 1450, 2000, 256, 5
    n = h * w
    if L % world == 0:
        for variant in ("default", "sharp_textured"):
            feats = torch.from_numpy(U.features(h, w, variant, seed=7)).to(dev)   # every rank the same guide
            mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).to(dev)
            l0, cnt = ls.label_shard(L, world, rank)
            g = torch.Generator(device=dev).manual_seed(100)
            e0 = torch.randn((n, L), generator=g, device=dev).abs_().mul_(4)[:, l0:l0 + cnt].contiguous()
            state = {}
            for mode in ("nccl_all_gather", "peer_memory_reads", "nccl_transpose_all_to_all",
                         "peer_memory_transpose"):
                box = {}

                def step():
                    lat = pkg.PermutohedralLattice(feats)
                    if mode == "nccl_all_gather":
                        res = ls.LabelShardedMeanField(ls.CudaBackend(lat)).run(e0, mu, niters, 1.0)
                    elif mode == "peer_memory_reads":
                        if "b" not in box:
                            box["b"] = ls.PeerMemoryBackend(lat, n, cnt)
                        box["b"].lat = lat
                        res = box["b"].run(e0, mu, niters, 1.0).clone()
                    elif mode == "nccl_transpose_all_to_all":
                        res = ls.TransposeExchangeMeanField(ls.CudaBackend(lat)).run(e0, mu, niters, 1.0)
                    else:
                        if "b" not in box:
                            box["b"] = ls.PeerTransposeMeanField(n, cnt, dev)
                        res = box["b"].run(lat, e0, mu, niters, 1.0)
                    state["M"], state["out"] = lat.num_vertices, res
                    lat.close()

                ms, clocks, err = None, None, None
                if rank == 0:
                    import sys
                    print(f"[bench_extra] config 4 {variant} x{world}: {mode}", file=sys.stderr, flush=True)
                try:
                    sampler = U.ClockSampler(local_rank).start() if rank == 0 else None
                    ms = timed(step)
                    clocks = sampler.stop() if sampler is not None else None
                except Exception as exc:     # a mode that cannot run on this box is reported, not fatal
                    err = repr(exc)[:200]
                if "b" in box:
                    try:
                        box["b"].close()
                    except Exception:
                        pass
                if rank == 0:
                    M = state.get("M", 0)
                    nb = U.bytes_frame(n, M, L, niters)
                    per_iter = {"nccl_all_gather": 4 * n * L * (world - 1) / world,
                                "peer_memory_reads": 4 * n * L * (world - 1) / world}.get(
                                    mode, 2 * 4 * n * L * (world - 1) / world / world)
                    e = {"workload": f"crf_{w}x{h}_L{L}_{niters}mf_{variant}_label_sharded",
                         "metric": "stereo_crf_frames_per_s_5mf", "unit": "frames/s", "n_gpus": world,
                         "scaling": "strong", "exchange": mode, "features": variant, "vertices": M,
                         "M_over_n": M / n, "clocks": clocks,
                         "nvlink_bytes_received_per_rank_per_iteration": per_iter}
                    if ms is not None:
                        e.update({"value": 1e3 / ms, "ms_per_frame": ms,
                                  "roofline": {"bound": "hbm", "scope": "whole frame over ALL ranks' HBM",
                                               "algorithmic_bytes": nb, "achieved": nb / (ms * 1e-3) / 1e9,
                                               "peak": peak * world, "unit": "GB/s",
                                               "frac": nb / (ms * 1e-3) / 1e9 / (peak * world),
                                               "peak_source": how}})
                    else:
                        e["unavailable"] = err
                    out.append(e)
                barrier()
            del e0, feats
            torch.cuda.empty_cache()

    # ---- config 5: 1920x1080 x 128, 8 frames per rank, no collective ----
    h, w, L, niters, frames = 1080, 1920, 128, 5, 8
    n = h * w
    feats = torch.from_numpy(U.features(h, w, "default", seed=7 + rank)).to(dev)
    mu = torch.from_numpy(synthetic.charb_matrix(L, 0.05)).to(dev)
    g = torch.Generator(device=dev).manual_seed(200 + rank)
    e0 = torch.randn((n, L), generator=g, device=dev).abs_().mul_(4)
    state = {}

    def step5():
        for _ in range(frames):
            lat = pkg.PermutohedralLattice(feats)
            state["out"] = lat.mean_field(e0, mu, niters, 1.0)
            state["M"] = lat.num_vertices
            lat.close()

    sampler = U.ClockSampler(local_rank).start() if rank == 0 else None
    ms = timed(step5)
    clocks = sampler.stop() if sampler is not None else None
    if rank == 0:
        M = state["M"]
        nb = U.bytes_frame(n, M, L, niters) * frames
        out.append({"workload": f"crf_{w}x{h}_L{L}_{niters}mf_default_batch{frames * world}_image_parallel",
                    "metric": "stereo_crf_frames_per_s_5mf", "unit": "frames/s", "n_gpus": world,
                    "scaling": "weak", "value": frames * world * 1e3 / ms, "ms_per_step": ms,
                    "frames_per_gpu": frames, "vertices": M, "M_over_n": M / n, "clocks": clocks,
                    "roofline": {"bound": "hbm", "scope": "whole frames, per GPU", "algorithmic_bytes": nb,
                                 "achieved": nb / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                 "frac": nb / (ms * 1e-3) / 1e9 / peak, "peak_source": how}})
    return out

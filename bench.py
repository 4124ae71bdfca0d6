"""bench.py -- stereo dense-CRF frames/s (5 mean-field iterations) on the lattice hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1]): Middlebury quarter-res 740x500 stereo pair,
64 disparity labels, 5-D (i, j, r, g, b) guide with s_ij = s_rgb = 0.1, charb
compatibility, 5 mean-field iterations, synthetic data.  A step is one frame
per rank: build the lattice from the guide features, then 5 fused iterations.
Image-parallel over ranks, no collective on the data path (weak scaling).

`value`  : frames/s with features and unaries already resident in HBM.
`e2e`    : frames/s through the module API (CRFasRNN.forward -> soft-argmax
           depth) from pinned HOST buffers, H2D of guide+logits and D2H of the
           depth map inside the timed region.
`roofline`: dominant kernel's algorithmic bytes / its mean launch duration
           (CUDA events on the launching stream, separate profiling pass right
           after the timed steps) against the measured HBM copy bandwidth.
`cpu_baseline`: the reference's CPU path (oracle/_ref when built, else the C
           restatement) on this box's host cores, rank 0, N = 1 only.
"""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, W, L, NITERS, D = 500, 740, 64, 5, 5
S_IJ = S_RGB = 0.1
GAMMA = 0.05
WORKLOAD = "middlebury_quarter_740x500_L64_5mf_forward"


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def _inputs(seed):
    import numpy as np
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    img = synthetic.image(H, W, seed)
    feats = synthetic.ijrgb_features(img, S_IJ, S_RGB)           # [n, 5]
    e0 = synthetic.unary(H, W, L, seed)                          # [n, L]
    mu = synthetic.charb_matrix(L, GAMMA)                        # [L, L]
    return np.ascontiguousarray(img), feats, e0, mu


class _ClockSampler:
    """SM clock and throttle reasons DURING the timed region, polled through NVML from a thread
    (the timed region lasts tens of milliseconds: `nvidia-smi -lms` is too coarse for it)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap"}

    def __init__(self, gpu_index):
        self.gpu, self.thread, self.stop_flag = gpu_index, None, False
        self.sm, self.bits, self.max_mhz, self.power = [], 0, None, []

    def _run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.gpu
            if vis:
                try:
                    idx = int(vis.split(",")[self.gpu])
                except Exception:
                    idx = self.gpu
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            while not self.stop_flag:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                try:
                    self.bits |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    try:
                        self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                    except Exception:
                        pass
                try:
                    self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
                except Exception:
                    pass
                time.sleep(0.002)
        except Exception:
            pass

    def start(self):
        import threading
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def stop(self):
        self.stop_flag = True
        if self.thread is not None:
            self.thread.join(timeout=5)
        out = {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": len(self.sm)}
        if self.sm:
            out["sm_mhz"] = statistics.median(self.sm)
        if self.power:
            out["power_w_max"] = max(self.power)
        out["reasons"] = sorted(name for bit, name in self.REASONS.items() if self.bits & bit)
        return out


def _algorithmic_bytes(kernel, n, M, C, launches_per_frame=None):
    """Per-launch algorithmic HBM bytes; d = 5, fp32."""
    d1 = D + 1
    if kernel.startswith(("mf_ws_kernel", "mf_row_kernel", "mf_tc_kernel", "mf_fused_kernel")):
        # fused iteration: first launch of a frame reads E0, writes P (+ splat record, lattice
        # rows written); the others also read P_prev, the slice record and the blurred lattice.
        first = 8 * n * C + 9 * n * d1 + 4 * M * C
        later = 12 * n * C + 15 * n * d1 + 8 * M * C
        k = NITERS if launches_per_frame is None else int(round(launches_per_frame))
        return (first + (k - 1) * later) / k
    if kernel.startswith("splat"):             # read src + entry lists, write lattice values
        return 4 * n * C + 9 * n * d1 + 4 * M * C
    if kernel.startswith("blur_kernel"):       # one axis: read + write values, neighbour ids
        return 8 * M * C + 8 * M
    if kernel.startswith("blur_"):             # all d+1 axes in one launch
        return d1 * (8 * M * C + 8 * M)
    if kernel.startswith("slice"):             # read values + entry lists + P + E0, write E
        return 4 * M * C + 6 * n * d1 + 12 * n * C
    if kernel.startswith("softmax_compat"):    # read E, write P, Mu once
        return 8 * n * C + 4 * C * C
    return None


def _ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, taken from the
    committed ncu --set full summary (profiles/ncu_traffic.json); None if not captured."""
    try:
        tab = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        for key, val in tab.items():
            if kernel.startswith(key):
                return val["dram_bytes_per_launch"]
    except Exception:
        pass
    return None


def cpu_frame_seconds(backend, feats, e0, mu, iters):
    """The reference's CPU path for `iters` mean-field iterations of one frame,
    pixel-major: crf_module.py:97-103 with W = filter - id (gaussian_matrix.py:352);
    the lattice is rebuilt inside every filter call, as the reference does."""
    import numpy as np
    import torch
    torch.set_num_threads(1)
    e0_t = torch.from_numpy(e0)
    mu_t = torch.from_numpy(mu)
    t0 = time.perf_counter()
    e = e0_t
    for _ in range(iters):
        p = torch.softmax(-e, dim=1) @ mu_t.t()
        f = torch.from_numpy(backend.filter(np.ascontiguousarray(p.numpy()), feats))
        e = e0_t + (f - p)
    return time.perf_counter() - t0


def _cpu_backend():
    from oracle.lattice_oracle import Oracle, Reference
    if Reference.available():
        try:
            return Reference()
        except Exception:
            pass
    return Oracle(faithful=True)


def _ref_worker(seed):
    img, feats, e0, mu = _inputs(seed)
    return cpu_frame_seconds(_cpu_backend(), feats, e0, mu, 1)


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on the host cores.
    The filter is single-threaded; its only parallelism is one process per image
    (gaussian_matrix.py:370-377), so `cores` frames run side by side.  Each step is a bounded
    sample: ONE mean-field iteration per frame (lattice rebuilt in the call, as the reference
    does), scaled x5 to the 5-iteration frame."""
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    procs = max(1, min(cores, 16))
    backend = _cpu_backend()
    ctx = mp.get_context("fork")
    times = []
    with ctx.Pool(procs) as pool:
        for step in range(args.warmup + args.steps):
            # workers time only the CRF iteration, not the synthetic-input generation
            dt = max(pool.map(_ref_worker, [1000 * step + i for i in range(procs)]))
            if step >= args.warmup:
                times.append(dt)
    per_step = sum(times) / len(times)           # procs frames x 1 iteration
    frame_s = per_step * NITERS                   # scaled to 5 iterations
    value = procs / frame_s
    sample = (f"{procs} frames in parallel processes, 1 of {NITERS} mean-field iterations each "
              f"(lattice rebuilt inside the filter call), time x{NITERS}")
    print(json.dumps({
        "impl": "reference", "metric": "stereo_crf_frames_per_s_5mf", "value": value,
        "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "h": H, "w": W, "labels": L, "mf_iters": NITERS, "d": D},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": procs, "kind": backend.kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def _bind_to_gpu_numa_node(local_rank):
    """One process per GPU: run (and first-touch the pinned e2e buffers) on the CPU cores next to the
    GPU, so that N ranks staging 100 MB per frame do not all pull from one socket's memory.  Best
    effort: returns the NUMA node or None."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local_rank)
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(path + "/numa_node").read())
        cpus = set()
        for part in open(path + "/local_cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        if cpus:
            os.sched_setaffinity(0, cpus & os.sched_getaffinity(0) or cpus)
        return node
    except Exception:
        return None


def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = _bind_to_gpu_numa_node(local_rank) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import depth_estimation_b200 as pkg
    cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
    mb = importlib.import_module("depth_estimation_b200.crf.mb_stereo_crf")
    N = pkg._native
    N.lib()

    img, feats, e0, mu = _inputs(seed=rank)
    n = H * W
    feats_d, e0_d, mu_d = (torch.from_numpy(x).to(dev) for x in (feats, e0, mu))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    state = {}

    def step():
        lat = pkg.PermutohedralLattice(feats_d)
        out = lat.mean_field(e0_d, mu_d, NITERS, 1.0)
        state["M"] = lat.num_vertices
        state["out"] = out
        lat.close()

    # ---- device-resident timing: per-step CUDA events, L2 flushed between steps ----
    for _ in range(args.warmup):
        step()
    barrier()
    sampler = _ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = N.launch_count()
    total_ms = 0.0
    for _ in range(args.steps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        a.record()
        step()
        b.record()
        torch.cuda.synchronize(dev)
        total_ms += a.elapsed_time(b)
    launches = N.launch_count() - launches0
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = reduce_max(total_ms)
    ms_per_step = total_ms / args.steps
    value = world * 1e3 / ms_per_step

    if args.only_step:
        if rank == 0:
            print(json.dumps({"metric": "stereo_crf_frames_per_s_5mf", "value": value,
                              "ms_per_step": ms_per_step, "only_step": True}))
        return

    # ---- end to end through the module API from pinned host buffers ----
    crf = cm.CRFasRNN(cm.charb(GAMMA), niters=NITERS).to(dev)
    guide_h = torch.from_numpy(feats).reshape(H, W, D).permute(2, 0, 1)[None].contiguous().pin_memory()
    logits_h = (-torch.from_numpy(e0)).reshape(H, W, L).permute(2, 0, 1)[None].contiguous().pin_memory()
    depth_h = torch.empty((1, 1, H, W), dtype=torch.float32).pin_memory()

    def e2e_step():
        with torch.no_grad():
            g = guide_h.to(dev, non_blocking=True)
            lg = logits_h.to(dev, non_blocking=True)
            depth = crf.depth(g, lg)      # = logits2average_depth(crf(g, lg)), read-out fused
            depth_h.copy_(depth, non_blocking=True)
            torch.cuda.synchronize(dev)
        gm.clear_lattice_cache()

    gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
    for _ in range(args.warmup):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_serial_s = reduce_max(time.perf_counter() - t0)
    e2e_serial = world * args.steps / e2e_serial_s

    # Same work with two frames in flight: two host threads, each with its own copy / compute
    # streams and device buffers, take alternate frames, so the H2D copy of one frame overlaps
    # the CRF of the other (the module call blocks its thread while the lattice's vertex count
    # comes back).  Every frame's H2D and D2H is inside the timed region.
    import threading
    NT = 2
    streams = [(torch.cuda.Stream(dev), torch.cuda.Stream(dev)) for _ in range(NT)]
    g_d = [torch.empty_like(guide_h, device=dev) for _ in range(NT)]
    lg_d = [torch.empty_like(logits_h, device=dev) for _ in range(NT)]
    dh = [torch.empty((1, 1, H, W), dtype=torch.float32).pin_memory() for _ in range(NT)]

    def worker(t, frames, errors):
        try:
            torch.cuda.set_device(dev)
            copy_st, comp_st = streams[t]
            ev = torch.cuda.Event()
            with torch.no_grad():
                for _ in range(frames):
                    with torch.cuda.stream(copy_st):
                        g_d[t].copy_(guide_h, non_blocking=True)
                        lg_d[t].copy_(logits_h, non_blocking=True)
                        ev.record(copy_st)
                    ev.synchronize()               # inputs are on the device
                    with torch.cuda.stream(comp_st):
                        depth = crf.depth(g_d[t], lg_d[t])
                        dh[t].copy_(depth, non_blocking=True)
                    comp_st.synchronize()          # this frame's depth map is on the host
                    gm.forget(g_d[t])
        except Exception as exc:                   # surface worker failures in the main thread
            errors.append(exc)

    def e2e_pipelined(steps):
        errors = []
        share = [steps // NT + (1 if t < steps % NT else 0) for t in range(NT)]
        threads = [threading.Thread(target=worker, args=(t, share[t], errors)) for t in range(NT)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        torch.cuda.synchronize(dev)
        if errors:
            raise errors[0]

    e2e_pipelined(max(args.warmup, 6 * NT))   # both threads past their first-touch allocations
    barrier()
    t0 = time.perf_counter()
    e2e_pipelined(args.steps)
    barrier()
    e2e_s = reduce_max(time.perf_counter() - t0)
    e2e_value = world * args.steps / e2e_s
    gm.clear_lattice_cache()

    # ---- extra: the stereo notebooks' whole pipeline, two images in -> depth map out ----
    # (cost volume crf/depth.py:34-53 with max_disp = L, guide fused into the lattice build,
    # 5 mean-field iterations, soft-argmax fused into the last slice; H2D of both images and
    # D2H of the depth map inside the timed region)
    left_h = torch.from_numpy(np.ascontiguousarray(img)).pin_memory()
    right_h = torch.from_numpy(np.ascontiguousarray(np.roll(img, -5, axis=1))).pin_memory()
    pair_depth_h = torch.empty((H, W), dtype=torch.float32).pin_memory()

    def pair_step():
        with torch.no_grad():
            lt = left_h.to(dev, non_blocking=True)
            rt = right_h.to(dev, non_blocking=True)
            pair_depth_h.copy_(cm.stereo_pair_depth(lt, rt, niters=NITERS, gamma=GAMMA, max_disp=L),
                               non_blocking=True)
            torch.cuda.synchronize(dev)

    for _ in range(args.warmup):
        pair_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        pair_step()
    barrier()
    pair_value = world * args.steps / reduce_max(time.perf_counter() - t0)

    # ---- per-kernel durations for the roofline (separate pass, same inputs) ----
    roofline = None
    if rank == 0:
        N.profile(True)
        for _ in range(3):
            flush.zero_()
            step()
        torch.cuda.synchronize(dev)
        N.profile(False)
        rep = N.profile_report()
        peak, how = _peaks()
        M = state["M"]
        top = max(rep.items(), key=lambda kv: kv[1][1])
        name, (cnt, ms) = top
        nbytes = _algorithmic_bytes(name, n, M, L, cnt / 3)
        frame_ms = sum(v[1] for v in rep.values()) / 3
        kernels = {k: {"launches_per_frame": v[0] / 3, "ms_per_frame": v[1] / 3,
                       "share": v[1] / 3 / frame_ms} for k, v in rep.items()}
        if nbytes is not None:
            achieved = nbytes / (ms / cnt * 1e-3) / 1e9
            roofline = {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peak,
                        "unit": "GB/s", "frac": achieved / peak, "traffic": _ncu_traffic(name),
                        "peak_source": how, "avg_launch_us": ms / cnt * 1e3,
                        "algorithmic_bytes_per_launch": nbytes, "vertices": M, "kernels": kernels}
        else:
            roofline = {"bound": "hbm", "kernel": name, "achieved": None, "peak": peak,
                        "unit": "GB/s", "frac": None, "traffic": None, "peak_source": how,
                        "kernels": kernels}

    # ---- CPU baseline: rank 0, N = 1 only ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        backend = _cpu_backend()
        secs = cpu_frame_seconds(backend, feats, e0, mu, NITERS)
        cpu = {"value": 1.0 / secs, "unit": "frames/s", "cores": 1, "kind": backend.kind,
               "sample": f"1 full frame ({NITERS} iterations, lattice rebuilt per filter call), "
                         f"single thread, {secs:.2f} s"}

    if rank == 0:
        print(json.dumps({
            "metric": "stereo_crf_frames_per_s_5mf", "value": value, "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "h": H, "w": W, "labels": L, "mf_iters": NITERS,
                       "d": D, "frames_per_step_per_gpu": 1, "parallelism": f"image-parallel x{world}", "rank0_numa_node": numa,
                       "l2": "flushed between timed steps (256 MiB write)",
                       "step": "lattice build + 5 fused mean-field iterations, inputs in HBM"},
            "e2e": {"value": e2e_value, "unit": "frames/s",
                    "h2d_bytes_per_step": int(guide_h.numel() + logits_h.numel()) * 4,
                    "d2h_bytes_per_step": int(depth_h.numel()) * 4,
                    "mode": "two frames in flight (two host threads, own streams): H2D of one overlaps the CRF of the other",
                    "serial_value": e2e_serial},
            "gpu_launches": int(launches) * world,
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "extra": {"lattice_filter_mpix_labels_per_s_incl_build": n * L * NITERS / (ms_per_step * 1e-3) / 1e6,
                      "vertices": state["M"],
                      "stereo_pair_to_depth_frames_per_s": pair_value,
                      "stereo_pair_h2d_bytes_per_step": int(2 * left_h.numel() * 4),
                      "stereo_pair_note": "images -> pl_cost_volume -> fused guide + 5 mean-field "
                                          "iterations -> fused soft-argmax; one frame at a time"},
        }))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--only-step", action="store_true",
                    help="run warm-up + timed device-resident steps only (for ncu)")
    args = ap.parse_args()
    # The JSON line must be the only thing on stdout: libraries that printf to fd 1 (NCCL's
    # version banner under NCCL_DEBUG) go to stderr, python's print keeps the real stdout.
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()

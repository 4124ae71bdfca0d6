"""Shared pieces of bench.py / bench_extra.py: measured peaks, NVML clock sampling during a timed
region, the synthetic feature variants, SURVEY 8d's algorithmic byte counts, bounded CPU samples
of the reference's own filter.  Measurement infrastructure only; no product code lives here."""
import importlib
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

D = 5

# Feature variants every workload is timed on:
#   default        ijrgbGuide's module default s_ij = s_rgb = 0.1 (crf_module.py:107)
#   sharp          the reference's canonical bilateral setting sigma_p = .01, sigma_c = .125
#                  (gaussian_matrix.py:498-505), same image
#   sharp_textured the same scales on a heavily textured image (noise 0.4): M/n 0.05-0.5 at MP sizes
VARIANTS = {
    "default": {"s_ij": 0.1, "s_rgb": 0.1, "noise": 0.05},
    "sharp": {"s_ij": 0.01, "s_rgb": 0.125, "noise": 0.05},
    "sharp_textured": {"s_ij": 0.01, "s_rgb": 0.125, "noise": 0.4},
}


def peaks():
    """(HBM GB/s, source): the driver-measured copy bandwidth, else the profiling guide's fallback."""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


_IMAGES = {}


def image(h, w, seed, noise):
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    key = (h, w, seed, noise)
    if key not in _IMAGES:
        if len(_IMAGES) > 4:
            _IMAGES.clear()
        _IMAGES[key] = synthetic.image(h, w, seed, noise)
    return _IMAGES[key]


def features(h, w, variant, seed=0):
    """[n, 5] ijrgb features (crf_module.py:111-114 order) of the full-resolution synthetic image."""
    synthetic = importlib.import_module("depth_estimation_b200.synthetic")
    v = VARIANTS[variant]
    return synthetic.ijrgb_features(image(h, w, seed, v["noise"]), v["s_ij"], v["s_rgb"])


class ClockSampler:
    """SM clock and throttle reasons DURING a timed region, polled through NVML from a thread
    (timed regions last milliseconds to seconds: `nvidia-smi -lms` is too coarse for the short ones)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap"}

    def __init__(self, gpu_index):
        self.gpu, self.thread, self.stop_flag = gpu_index, None, False
        self.sm, self.bits, self.max_mhz, self.power = [], 0, None, []
        import threading
        self.ready = threading.Event()     # NVML initialised, sampling has begun

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
            self.ready.set()
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
        finally:
            self.ready.set()

    def start(self):
        """Returns once NVML is up and the first samples are being taken: a timed region of a few
        milliseconds must not be over before the sampler thread has finished initialising."""
        import threading
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        self.ready.wait(timeout=5.0)
        return self

    def stop(self):
        self.stop_flag = True
        if self.thread is not None:
            self.thread.join(timeout=5)
        out = {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": len(self.sm)}
        if self.sm:
            out["sm_mhz"] = statistics.median(self.sm)
            out["sm_mhz_min"] = min(self.sm)
        if self.power:
            out["power_w_max"] = max(self.power)
        out["reasons"] = sorted(name for bit, name in self.REASONS.items() if self.bits & bit)
        return out


# ---- algorithmic HBM bytes, SURVEY.md section 8(d), d = 5, fp32 -------------------------------------
def table_slots(n):
    """Slots of the build's hash table: pow2_ceil(1.25 E + 1024) (csrc/pl_build.cu)."""
    e = n * (D + 1)
    need, p = e + e // 4 + 1024, 1
    while p < need:
        p <<= 1
    return p


def bytes_apply(n, M, C):
    """Filter apply on a cached lattice: 8nC + 96n + 56MC + 48M."""
    return 8 * n * C + 96 * n + 56 * M * C + 48 * M


def bytes_build(n, M, from_rgb=False):
    """Lattice build: 4nd (12n from RGB) + 8n(d+1) + 2Md + 8M(d+1) + 8T."""
    return (12 * n if from_rgb else 4 * n * D) + 8 * n * (D + 1) + 2 * M * D + 8 * M * (D + 1) + 8 * table_slots(n)


def bytes_iter(n, M, L):
    """One mean-field iteration: apply(C = L) + read E0 + write E/Q."""
    return bytes_apply(n, M, L) + 8 * n * L


def bytes_frame(n, M, L, niters):
    return bytes_build(n, M) + niters * bytes_iter(n, M, L)


def kernel_bytes(kernel, n, M, C, niters, launches_per_frame=None):
    """Per-launch algorithmic bytes of one of OUR kernels, for the dominant-kernel
    roofline.  None: no formula for this kernel."""
    d1 = D + 1
    if kernel.startswith(("mf_ws_kernel", "mf_tc_kernel", "mf_fused_kernel", "mf_big")):
        first = 8 * n * C + 9 * n * d1 + 4 * M * C
        later = 12 * n * C + 15 * n * d1 + 8 * M * C
        k = niters if launches_per_frame is None else max(1, int(round(launches_per_frame)))
        return (first + (k - 1) * later) / k
    if kernel.startswith("splat"):
        return 4 * n * C + 9 * n * d1 + 4 * M * C
    if kernel.startswith("blur_kernel"):
        return 8 * M * C + 8 * M
    if kernel.startswith("blur_"):
        return d1 * (8 * M * C + 8 * M)
    if kernel.startswith("slice_kernel"):
        return 4 * M * C + 8 * n * d1 + 4 * n * C
    if kernel.startswith("slice"):     # with the E0 + (KP - P) epilogue inside a CRF frame, plain in a filter
        return 4 * M * C + 6 * n * d1 + (12 if niters > 0 else 4) * n * C
    if kernel.startswith(("softmax_compat", "compat_tc")):
        return 8 * n * C + 4 * C * C
    return None


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed
    `ncu --set full` summary (profiles/ncu_traffic.json); None if not captured."""
    try:
        tab = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        for key, val in tab.items():
            if kernel.startswith(key):
                return val["dram_bytes_per_launch"]
    except Exception:
        pass
    return None


# ---- the reference's CPU path, bounded samples ------------------------------------------------------
def cpu_backend():
    from oracle.lattice_oracle import Oracle, Reference
    if Reference.available():
        try:
            return Reference()
        except Exception:
            pass
    return Oracle(faithful=True)


def cpu_frame_seconds(backend, feats, e0, mu, iters):
    """The reference's CPU path for `iters` mean-field iterations of one frame, pixel-major:
    crf_module.py:97-103 with W = filter - id (gaussian_matrix.py:352); the lattice is rebuilt inside
    every filter call, as the reference does."""
    import numpy as np
    import torch
    torch.set_num_threads(1)
    e0_t, mu_t = torch.from_numpy(e0), torch.from_numpy(mu)
    t0 = time.perf_counter()
    e = e0_t
    for _ in range(iters):
        p = torch.softmax(-e, dim=1) @ mu_t.t()
        f = torch.from_numpy(backend.filter(np.ascontiguousarray(p.numpy()), feats))
        e = e0_t + (f - p)
    return time.perf_counter() - t0


class CpuFilterModel:
    """Bounded CPU sample of the reference filter on one feature matrix: the first `rows` points
    are filtered with 1 and with 16 value channels (one thread), giving t(n, C) ~ n (a + b C); larger
    shapes are EXTRAPOLATED from it (and labelled so)."""

    def __init__(self, backend, feats, rows=98304):
        import numpy as np
        self.kind = backend.kind
        r = min(rows, feats.shape[0])
        f = np.ascontiguousarray(feats[:r])
        rng = np.random.default_rng(0)
        t = {}
        for c in (1, 16):
            src = rng.random((r, c)).astype(np.float32)
            t0 = time.perf_counter()
            backend.filter(src, f)
            t[c] = (time.perf_counter() - t0) / r
        self.b = max((t[16] - t[1]) / 15.0, 0.0)
        self.a = max(t[1] - self.b, 0.0)
        self.rows, self.spent = r, (t[1] + t[16]) * r

    def filter_seconds(self, n, C):
        return n * (self.a + self.b * C)

    def sample_text(self):
        return (f"first {self.rows} points filtered with 1 and 16 channels on one thread "
                f"({self.spent:.2f} s), t = n(a + bC) extrapolated")


def cpu_compat_seconds_per_row(L):
    """softmax + [L x L] compatibility product per pixel row on one thread (bounded sample)."""
    import torch
    torch.set_num_threads(1)
    rows = max(2048, min(65536, (1 << 24) // (L * L)))
    e = torch.rand((rows, L))
    mu = torch.rand((L, L))
    t0 = time.perf_counter()
    _ = torch.softmax(-e, dim=1) @ mu.t()
    return (time.perf_counter() - t0) / rows

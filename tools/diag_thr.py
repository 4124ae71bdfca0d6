import importlib, sys, time, os, threading
sys.path.insert(0, os.getcwd())
import torch
import bench
import depth_estimation_b200 as pkg
cm = importlib.import_module("depth_estimation_b200.crf.crf_module")
mb = importlib.import_module("depth_estimation_b200.crf.mb_stereo_crf")
gm = importlib.import_module("depth_estimation_b200.crf.gaussian_matrix")
dev = torch.device("cuda", 0)
H, W, L, D = bench.H, bench.W, bench.L, bench.D
img, feats, e0, mu = bench._inputs(0)
crf = cm.CRFasRNN(cm.charb(bench.GAMMA), niters=5).to(dev)
guide_h = torch.from_numpy(feats).reshape(H, W, D).permute(2, 0, 1)[None].contiguous().pin_memory()
logits_h = (-torch.from_numpy(e0)).reshape(H, W, L).permute(2, 0, 1)[None].contiguous().pin_memory()
NT = 2
streams = [(torch.cuda.Stream(dev), torch.cuda.Stream(dev)) for _ in range(NT)]
g_d = [torch.empty_like(guide_h, device=dev) for _ in range(NT)]
lg_d = [torch.empty_like(logits_h, device=dev) for _ in range(NT)]
dh = [torch.empty((1, 1, H, W), dtype=torch.float32).pin_memory() for _ in range(NT)]
T0 = time.perf_counter()
log = [[] for _ in range(NT)]
def worker(t, frames):
    torch.cuda.set_device(dev)
    copy_st, comp_st = streams[t]
    ev = torch.cuda.Event()
    with torch.no_grad():
        for _ in range(frames):
            a = time.perf_counter() - T0
            with torch.cuda.stream(copy_st):
                g_d[t].copy_(guide_h, non_blocking=True)
                lg_d[t].copy_(logits_h, non_blocking=True)
                ev.record(copy_st)
            b = time.perf_counter() - T0
            ev.synchronize()
            b2 = time.perf_counter() - T0
            with torch.cuda.stream(comp_st):
                comp_st.wait_event(ev)
                depth = mb.logits2average_depth(crf(g_d[t], lg_d[t]))
                dh[t].copy_(depth, non_blocking=True)
            c = time.perf_counter() - T0
            comp_st.synchronize()
            d = time.perf_counter() - T0
            gm.forget(g_d[t])
            log[t].append((a, b, b2, c, d))
for rep in range(2):
    T0 = time.perf_counter(); log = [[] for _ in range(NT)]
    ths = [threading.Thread(target=worker, args=(t, 6)) for t in range(NT)]
    for th in ths: th.start()
    for th in ths: th.join()
    tot = time.perf_counter() - T0
    print("ms/frame", tot / 12 * 1e3)
for t in range(NT):
    for (a, b, b2, c, d) in log[t]:
        print(f"thr{t}: start {a*1e3:7.2f} copy-enq {b*1e3:7.2f} copy-done {b2*1e3:7.2f} crf-ret {c*1e3:7.2f} done {d*1e3:7.2f}")

import importlib, sys, time, os
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
copy_st, comp_st = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
g_d = [torch.empty_like(guide_h, device=dev) for _ in range(2)]
lg_d = [torch.empty_like(logits_h, device=dev) for _ in range(2)]
dh = [torch.empty((1, 1, H, W), dtype=torch.float32).pin_memory() for _ in range(2)]
A = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)

def run(kind, steps=12):
    evs = []
    ev_in = [None, None]; ev_out = [None, None]
    base = torch.cuda.Event(enable_timing=True); base.record(); 
    cpu = []
    t0 = time.perf_counter()
    for i in range(steps):
        k = i & 1
        with torch.cuda.stream(copy_st):
            if i >= 2: copy_st.wait_event(ev_out[k])
            s = torch.cuda.Event(enable_timing=True); s.record(copy_st)
            g_d[k].copy_(guide_h, non_blocking=True)
            lg_d[k].copy_(logits_h, non_blocking=True)
            e = torch.cuda.Event(enable_timing=True); e.record(copy_st)
            ev_in[k] = e
        c0 = time.perf_counter()
        with torch.cuda.stream(comp_st), torch.no_grad():
            comp_st.wait_event(ev_in[k])
            s2 = torch.cuda.Event(enable_timing=True); s2.record(comp_st)
            if kind == "crf":
                depth = mb.logits2average_depth(crf(g_d[k], lg_d[k]))
                dh[k].copy_(depth, non_blocking=True)
            else:
                for _ in range(2): B = A @ A
            e2 = torch.cuda.Event(enable_timing=True); e2.record(comp_st)
            ev_out[k] = e2
        c1 = time.perf_counter()
        evs.append((s, e, s2, e2)); cpu.append((c0 - t0, c1 - t0))
    torch.cuda.synchronize()
    total = time.perf_counter() - t0
    print(kind, "ms/frame", total / steps * 1e3)
    for i, (s, e, s2, e2) in enumerate(evs):
        print(f"  {i}: copy [{base.elapsed_time(s):7.2f},{base.elapsed_time(e):7.2f}] comp [{base.elapsed_time(s2):7.2f},{base.elapsed_time(e2):7.2f}] cpu call [{cpu[i][0]*1e3:7.2f},{cpu[i][1]*1e3:7.2f}]")
    gm.clear_lattice_cache()

gm.set_lattice_cache_limit(6)
run("mm"); run("mm")
run("crf"); run("crf")

"""Markdown tables for DESIGN.md section 7 / BASELINE.md section 2d from the committed bench lines.

    python tools/make_tables.py [profiles/r2f]      # prefix of <prefix>_bench.json, _bench_reference.json, _bench_{2,4,8}gpu.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pre = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r2a")


def load(name):
    for cand in (f"{pre}_{name}.json", os.path.join(ROOT, "profiles", f"r2k_{name}.json"), os.path.join(ROOT, "profiles", f"r2h_{name}.json"),
                 os.path.join(ROOT, "profiles", f"r2a_{name}.json")):
        if os.path.exists(cand):
            return json.load(open(cand)), os.path.basename(cand)
    return None, None


one, one_name = load("bench")
ref, _ = load("bench_reference")
multi = {g: load(f"bench_{g}gpu") for g in (2, 4, 8)}

print(f"Source: `profiles/{one_name}`" + "".join(f", `profiles/{n}`" for _, n in multi.values() if n) + ".\n")
r = one
print("| headline (740×500 × 64, 5 iterations, default scales) | value |")
print("|---|---|")
print(f"| device-resident, 20 steps, L2 flushed | {r['ms_per_step']:.3f} ms/frame, **{r['value']:.0f} frames/s** (clocks {r['clocks']['sm_mhz']:.0f} MHz, reasons {r['clocks']['reasons']}) |")
su = r["sustained"]
print(f"| sustained, {su['steps']} frames back to back ({su['seconds']:.2f} s) | {su['ms_per_step']:.3f} ms/frame, {su['value']:.0f} frames/s (median {su['clocks']['sm_mhz']:.0f} MHz, min {su['clocks'].get('sm_mhz_min')}, {su['clocks'].get('power_w_max', 0):.0f} W, reasons {su['clocks']['reasons']}) |")
e = r["e2e"]
print(f"| e2e, images in → depth out (`stereo_pair_depth`, {e['h2d_bytes_per_step'] / 1e6:.1f} MB H2D + {e['d2h_bytes_per_step'] / 1e6:.1f} MB D2H per frame) | **{e['value']:.0f} frames/s** two frames in flight, {e['serial_value']:.0f} one at a time |")
m = e["module_call"]
print(f"| e2e, module call (`CRFasRNN.depth`, {m['h2d_bytes_per_step'] / 1e6:.0f} MB H2D per frame: PCIe bound) | {m['value']:.0f} frames/s pipelined, {m['serial_value']:.0f} serial |")
if ref:
    print(f"| reference CPU arm (`oracle/_ref`, {ref['cpu_baseline']['cores']} processes, full frames) | {ref['value']:.2f} frames/s → e2e ratio {e['value'] / ref['value']:.0f}×, device ratio {r['value'] / ref['value']:.0f}× |")
if r.get("cpu_baseline"):
    print(f"| reference CPU, one core | {r['cpu_baseline']['value']:.3f} frames/s |")
ro = r["roofline"]
print(f"| dominant kernel `{ro['kernel']}` | {ro['avg_launch_us']:.1f} µs per launch, {ro['achieved']:.0f} GB/s algorithmic = **{ro['frac']:.3f}** of {ro['peak']:.0f} GB/s; ncu DRAM traffic {ro['traffic'] / 1e6:.1f} MB vs {ro['algorithmic_bytes_per_launch'] / 1e6:.1f} MB algorithmic |")
print(f"| whole frame on SURVEY 8d bytes | {ro['frame']['algorithmic_bytes'] / 1e9:.2f} GB → {ro['frame']['frac']:.3f} of peak |")
shares = sorted(ro["kernels"].items(), key=lambda kv: -kv[1]["ms_per_frame"])
print("| per-frame shares | " + ", ".join(f"{k.split('<')[0]} {v['share'] * 100:.1f} %" for k, v in shares[:8]) + " |")

print("\n| workload | features | M/n | ours 1×B200 | roofline frac (SURVEY 8d bytes) | dominant kernel (its frac) | reference CPU, 1 core (extrapolated) | SM MHz |")
print("|---|---|---|---|---|---|---|---|")
train = [c for c in (r["extra"]["configs"] or []) if "train_step_ms" in c]
for c in r["extra"]["configs"] or []:
    if "train_step_ms" in c:
        continue
    dk = c["roofline"].get("dominant_kernel") or {}
    val = (f"{c['value']:.1f} frames/s ({c['ms_per_frame']:.2f} ms)" if c["unit"] == "frames/s"
           else f"{c['value'] / 1e3:.1f} Gpix·labels/s apply ({c['ms_apply']:.3f} ms), {c['value_incl_build'] / 1e3:.1f} incl. build ({c['ms_build']:.2f} ms)")
    cpu = c.get("cpu_baseline", {})
    cpuv = (f"{cpu['value']:.4g} {cpu['unit']}" if cpu else "–")
    name = c["workload"].rsplit("_" + c["features"], 1)[0]
    print(f"| {name} | {c['features']} | {c['M_over_n']:.3f} | {val} | {c['roofline']['frac']:.3f} | "
          f"{dk.get('kernel', '').split('<')[0]} ({dk.get('frac', 0) or 0:.2f}) | {cpuv} | {c['clocks']['sm_mhz']:.0f} |")

if train:
    print("\nTraining step (config 2: 740×500 × 64, 5 iterations, forward + backward): " + "; ".join(
        f"{c['train_step_ms']:.2f} ms with grads wrt {c['grads']} ({c['clocks']['sm_mhz']:.0f} MHz)" for c in train) + ".")
print("\n| GPUs | image-parallel 740×500×64: device frames/s (×1 GPU) | e2e frames/s (×) | 1080p × 128 batch, frames/s | config 4 label-sharded, ms/frame: all-gather / peer reads / transpose a2a / transpose peer kernel |")
print("|---|---|---|---|---|")
base_v, base_e = r["value"], r["e2e"]["value"]
c5 = [c for c in r["extra"]["configs"] if c["workload"].startswith("crf_1920x1080_L128") and c["features"] == "default"]
c4 = {c["features"]: c for c in r["extra"]["configs"] if c["workload"].startswith("crf_2000x1450_L256")}
print(f"| 1 | {base_v:.0f} | {base_e:.0f} | {c5[0]['value']:.1f} | one GPU, unsharded: {c4['default']['ms_per_frame']:.1f} (default), {c4['sharp_textured']['ms_per_frame']:.1f} (sharp textured) |")
for g, (m, mname) in multi.items():
    if not m:
        continue
    # a row from an older run set is scaled against THAT set's own 1-GPU line, and says so
    tag = mname.split("_bench")[0]
    note = ""
    bv, be = base_v, base_e
    if tag != one_name.split("_bench")[0]:
        older = os.path.join(ROOT, "profiles", f"{tag}_bench.json")
        if os.path.exists(older):
            o = json.load(open(older))
            bv, be = o["value"], o["e2e"]["value"]
        note = f" [{tag}: measured before this round's last optimisations; × against {tag}'s own 1-GPU line]"
    cf = m["extra"]["configs"] or []
    ls = {}
    for c in cf:
        if "label_sharded" in c["workload"]:
            ls.setdefault(c["features"], []).append(f"{c['ms_per_frame']:.1f}" if "ms_per_frame" in c else "n/a")
    ip = [c for c in cf if "image_parallel" in c["workload"]]
    print(f"| {g}{note} | {m['value']:.0f} ({m['value'] / bv:.2f}×) | {m['e2e']['value']:.0f} ({m['e2e']['value'] / be:.2f}×) | "
          f"{ip[0]['value']:.1f} | default: {' / '.join(ls.get('default', []))}; sharp textured: {' / '.join(ls.get('sharp_textured', []))} |")

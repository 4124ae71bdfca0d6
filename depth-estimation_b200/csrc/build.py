"""Build libpl_lattice.so in-tree for sm_100a (nvcc cross-compiles without a GPU).

    python depth-estimation_b200/csrc/build.py [--force] [--verbose]

Output: depth-estimation_b200/libpl_lattice.so (git-ignored; travels to the GPU
box with the gpurun snapshot).  No torch headers are involved: the library is a
plain C ABI (include/pl_lattice.h) that the Python host side loads with ctypes.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
OUT = os.path.join(PKG, "libpl_lattice.so")
SOURCES = ["pl_api.cu", "pl_build.cu", "pl_filter.cu", "pl_meanfield.cu", "pl_tiles.cu", "pl_fused.cu", "pl_fused_tc.cu", "pl_fused_ws.cu", "pl_cost.cu", "pl_readout.cu", "pl_compat_tc.cu", "pl_backward.cu", "pl_noalloc.cu"]
HEADERS = ["pl_common.cuh", "pl_tc.cuh", os.path.join("..", "..", "include", "pl_lattice.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-extended-lambda"]


def _stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(HERE, f)) > t for f in SOURCES + HEADERS + ["build.py"])


def build(force=False, verbose=False):
    if not force and not _stale():
        return OUT
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(HERE, src.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", os.path.join(HERE, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- {src}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    subprocess.check_call([NVCC, "-shared", "-o", OUT] + objs + ["-lcudart"])
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))

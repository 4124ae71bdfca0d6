"""Build oracle/_ref/libref_lattice.so from the reference sources IN PLACE.

TEST INFRASTRUCTURE.  Compiles oracle/ref_driver.cpp, which #includes
/root/reference/crf/lattice/lite/permutohedral.h by absolute path (nothing is
copied), with the flags torch's extension builder would use (-O2, C++17, no
-march) -- i.e. the way crf/gaussian_matrix.py:15 builds the reference -- and
links against libtorch.  Output: oracle/_ref/libref_lattice.so (git-ignored,
travels to the GPU box with the snapshot).  No-op when /root/reference is
absent (the GPU box): the prebuilt file is used as is.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_HEADER = "/root/reference/crf/lattice/lite/permutohedral.h"
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "libref_lattice.so")


def build(force=False):
    if not os.path.exists(REF_HEADER):
        return OUT if os.path.exists(OUT) else None
    src = os.path.join(HERE, "ref_driver.cpp")
    if (not force and os.path.exists(OUT)
            and os.path.getmtime(OUT) >= max(os.path.getmtime(src), os.path.getmtime(REF_HEADER))):
        return OUT
    from torch.utils import cpp_extension as ce
    import torch
    os.makedirs(OUT_DIR, exist_ok=True)
    inc = []
    for p in ce.include_paths():
        inc += ["-isystem", p]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = (["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w",
            f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
            f'-DREF_HEADER="{REF_HEADER}"'] + inc +
           [src, "-o", OUT, f"-L{libdir}", f"-Wl,-rpath,{libdir}",
            "-ltorch", "-ltorch_cpu", "-lc10"])
    subprocess.check_call(cmd)
    return OUT


CUDA_SRC = "/root/reference/crf/lattice/cuda_lattice/permutohedral.cu"
OUT_CUDA = os.path.join(OUT_DIR, "libref_cuda_lattice.so")


def build_cuda(force=False):
    """The reference's stand-alone CUDA lattice (crf/lattice/cuda_lattice/permutohedral.cu, which
    #includes hash_table.cu and the cutil.h vendored beside it) compiled IN PLACE for sm_100a in
    LIBRARY mode: exports `filter(float* im, float* ref, int pd, int vd, int w, int h, bool
    accurate)` (:523-529), host pointers in and out, vd <= 3.  A second SPEED baseline for the
    filter sweep (bench_configs.py --workload ref_cuda_filter); its arithmetic is Adams'
    normalised Gaussian, not the lite path's, so it is never used for parity."""
    if not os.path.exists(CUDA_SRC):
        return OUT_CUDA if os.path.exists(OUT_CUDA) else None
    if not force and os.path.exists(OUT_CUDA) and os.path.getmtime(OUT_CUDA) >= os.path.getmtime(CUDA_SRC):
        return OUT_CUDA
    os.makedirs(OUT_DIR, exist_ok=True)
    cmd = ["nvcc", "-w", "-O2", "-DLIBRARY", "-gencode", "arch=compute_100a,code=sm_100a", "-shared",
           "-Xcompiler", "-fPIC", "-Xcompiler", "-w", "-I", os.path.dirname(CUDA_SRC),
           "-include", os.path.join(HERE, "cuda_symbol_compat.h"), "-o", OUT_CUDA, CUDA_SRC, os.path.join(HERE, "cutil_shim.c")]
    subprocess.check_call(cmd, cwd=OUT_DIR)
    return OUT_CUDA


if __name__ == "__main__":
    out = build(force="--force" in sys.argv)
    print(out if out else "reference sources absent and no prebuilt oracle/_ref")
    out = build_cuda(force="--force" in sys.argv)
    print(out if out else "reference CUDA sources absent and no prebuilt oracle/_ref")

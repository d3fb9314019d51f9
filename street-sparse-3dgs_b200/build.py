#!/usr/bin/env python
"""Build libh3dgs_b200.so (the C-ABI library of include/h3dgs_b200.h) for sm_100a with nvcc.

In-tree output: street-sparse-3dgs_b200/lib/libh3dgs_b200.so (git-ignored, shipped to the GPU box).
Floating-point flags are the reference's (nvcc defaults: -fmad=true, IEEE div/sqrt, no fast-math):
sort keys and tile rectangles must come out bit-identical.
"""
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "lib")
OBJ_DIR = os.path.join(OUT_DIR, "obj")
LIB = os.path.join(OUT_DIR, "libh3dgs_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "-Xptxas", "-v", "--expt-relaxed-constexpr"]
SOURCES = ["api.cu", "preprocess_fwd.cu", "scan.cu", "binning.cu", "radix_sort.cu", "blend_fwd.cu",
           "blend_bwd.cu", "blend_bwd2.cu", "preprocess_bwd.cu", "cut.cu", "sparse_adam.cu", "loss.cu", "peaks.cu", "hier_build.cu", "dp_exchange.cu", "densify.cu",
           "hier_io.cpp"]


def _digest():
    h = hashlib.sha256()
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in sorted(os.listdir(root)):
            with open(os.path.join(root, f), "rb") as fh:
                h.update(f.encode() + b"\0" + fh.read())
    h.update(" ".join(ARCH + COMMON).encode())
    return h.hexdigest()


def _compile(src):
    obj = os.path.join(OBJ_DIR, src + ".o")
    cmd = [NVCC] + ARCH + COMMON + ["-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(obj + ".log", "w") as fh:
        fh.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force=False, verbose=False):
    os.makedirs(OBJ_DIR, exist_ok=True)
    stamp = os.path.join(OUT_DIR, "build.sha256")
    dg = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read().strip() == dg:
        return LIB
    with cf.ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(_compile, SOURCES))
    cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(stamp, "w") as fh:
        fh.write(dg)
    if verbose:
        for s in SOURCES:
            print(open(os.path.join(OBJ_DIR, s + ".o.log")).read())
    return LIB


EXT_NAME = "_h3dgs_torch"
EXT = os.path.join(OUT_DIR, EXT_NAME + ".so")


def build_torch_ext(force=False, verbose=False):
    """The torch extension (csrc/torch_binding.cpp -> lib/_h3dgs_torch.so): host C++ over the C ABI, linked against
    lib/libh3dgs_b200.so through an $ORIGIN rpath.  In-tree output, like the library itself."""
    import shutil
    build()
    src = os.path.join(CSRC, "torch_binding.cpp")
    stamp = os.path.join(OUT_DIR, "ext.sha256")
    h = hashlib.sha256()
    for f in (src, os.path.join(HERE, "..", "include", "h3dgs_b200.h")):
        h.update(open(f, "rb").read())
    import torch
    h.update(torch.__version__.encode())
    dg = h.hexdigest()
    if not force and os.path.exists(EXT) and os.path.exists(stamp) and open(stamp).read().strip() == dg:
        return EXT
    import subprocess
    import sysconfig
    from torch.utils.cpp_extension import include_paths, library_paths
    inc = include_paths() + [sysconfig.get_paths()["include"], "/usr/local/cuda/include",
                             os.path.join(HERE, "..", "include")]
    cmd = (["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-Wno-interference-size", "-DTORCH_EXTENSION_NAME=" + EXT_NAME,
            "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=" + str(int(torch._C._GLIBCXX_USE_CXX11_ABI))]
           + ["-I" + d for d in inc] + [src, "-o", EXT, "-L" + OUT_DIR, "-lh3dgs_b200", "-Wl,-rpath,$ORIGIN"]
           + ["-L" + d for d in library_paths()] + ["-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch", "-ltorch_python",
                                                     "-L/usr/local/cuda/lib64", "-lcudart"])
    if verbose:
        print(" ".join(cmd))
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as fh:
        fh.write(dg)
    return EXT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_torch_ext(force="--force" in sys.argv, verbose="-v" in sys.argv))

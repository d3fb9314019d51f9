#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY -- builds the *unmodified* reference CUDA extensions for sm_100a.

The reference sources are compiled where they lie under /root/reference (never copied);
only build products land in ``oracle/_ref/`` (git-ignored, but shipped to the GPU box):

    oracle/_ref/ref_hr_C.so   <- submodules/hierarchy-rasterizer  (rasterize_gaussians[_backward])
    oracle/_ref/ref_gh_C.so   <- submodules/gaussianhierarchy     (expand_to_size, get_interpolation_weights, ...)
    oracle/_ref/ref_switch_tool <- submodules/gaussianhierarchy/runtime_switching.cu under oracle/reftool/ref_switch_tool.cu
                                 (Switching::changeToSizeStep, the viewer's incremental cut) -- nvcc, no torch
    oracle/_ref/ref_gh_tool   <- submodules/gaussianhierarchy host C++ (HierarchyWriter / HierarchyLoader, the
                                 hierarchy Creator's CPU stages) under the command-line driver
                                 oracle/reftool/ref_gh_tool.cpp -- plain g++, no torch

These are the bit-exact oracle for ``tests/ -m gpu`` and the timed arm of
``bench.py --impl reference``.  The product path never imports them.

The reference's own build system (setup.py / cmake) is not run.  The one deviation from a
stock build is ``-include cstdint``: cuda_rasterizer/rasterizer_impl.h uses std::uintptr_t
without the header, which gcc 13 rejects.
"""
import os
import sys
import shutil

REF = os.environ.get("H3DGS_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")


TOOL_SOURCES = ("hierarchy_writer.cpp", "hierarchy_loader.cpp", "PointbasedKdTreeGenerator.cpp", "ClusterMerger.cpp",
                "rotation_aligner.cpp", "writer.cpp")


def build_tool(verbose=True):
    """g++ on the reference's own host sources where they lie + our driver; output oracle/_ref/ref_gh_tool."""
    import hashlib
    import subprocess
    gh = os.path.join(REF, "submodules/gaussianhierarchy")
    driver = os.path.join(HERE, "reftool", "ref_gh_tool.cpp")
    out = os.path.join(OUT, "ref_gh_tool")
    stamp = out + ".sha256"
    h = hashlib.sha256(open(driver, "rb").read() + " ".join(TOOL_SOURCES).encode()).hexdigest()
    if os.path.exists(out) and os.path.exists(stamp) and open(stamp).read().strip() == h:
        return True
    os.makedirs(OUT, exist_ok=True)
    cmd = ["g++", "-O2", "-std=c++17", "-w", "-fopenmp", "-I" + gh, "-I" + os.path.join(gh, "dependencies/eigen"), driver] + \
        [os.path.join(gh, s) for s in TOOL_SOURCES] + ["-o", out]
    if verbose:
        print("[build_ref]", " ".join(cmd))
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as fh:
        fh.write(h)
    return True


def build_switch_tool(verbose=True):
    """nvcc on the reference's runtime_switching.cu where it lies + our driver; output oracle/_ref/ref_switch_tool
    (Switching::changeToSizeStep, the viewer's incremental cut, has no Python binding in the reference)."""
    import hashlib
    import subprocess
    gh = os.path.join(REF, "submodules/gaussianhierarchy")
    driver = os.path.join(HERE, "reftool", "ref_switch_tool.cu")
    out = os.path.join(OUT, "ref_switch_tool")
    stamp = out + ".sha256"
    h = hashlib.sha256(open(driver, "rb").read()).hexdigest()
    if os.path.exists(out) and os.path.exists(stamp) and open(stamp).read().strip() == h:
        return True
    os.makedirs(OUT, exist_ok=True)
    cmd = [os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc"), "-gencode", "arch=compute_100a,code=sm_100a", "-O2",
           "-std=c++17", "-w", "-I" + gh, "-I" + os.path.join(gh, "dependencies/eigen"), driver,
           os.path.join(gh, "runtime_switching.cu"), "-o", out]
    if verbose:
        print("[build_ref]", " ".join(cmd))
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as fh:
        fh.write(h)
    return True


def build(which=("hr", "gh", "tool"), verbose=True):
    if not os.path.isdir(REF):
        print(f"[build_ref] {REF} not present; nothing to do", file=sys.stderr)
        return False
    if "tool" in which:
        build_tool(verbose)
        build_switch_tool(verbose)
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0a"
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 4))
    from torch.utils.cpp_extension import load

    os.makedirs(OUT, exist_ok=True)
    ok = True
    if "hr" in which and not os.path.exists(os.path.join(OUT, "ref_hr_C.so")):
        hr = os.path.join(REF, "submodules/hierarchy-rasterizer")
        bdir = os.path.join(OUT, "build_hr")
        os.makedirs(bdir, exist_ok=True)
        load(
            name="ref_hr_C",
            sources=[os.path.join(hr, s) for s in (
                "cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
                "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp")],
            extra_include_paths=[os.path.join(hr, "third_party/glm")],
            extra_cflags=["-O3", "-include", "cstdint", "-w"],
            extra_cuda_cflags=["-include", "cstdint", "-w", "-Xcompiler", "-fno-gnu-unique"],
            build_directory=bdir, verbose=verbose, is_python_module=False)
        shutil.copy2(os.path.join(bdir, "ref_hr_C.so"), os.path.join(OUT, "ref_hr_C.so"))
    if "gh" in which and not os.path.exists(os.path.join(OUT, "ref_gh_C.so")):
        gh = os.path.join(REF, "submodules/gaussianhierarchy")
        bdir = os.path.join(OUT, "build_gh")
        os.makedirs(bdir, exist_ok=True)
        load(
            name="ref_gh_C",
            sources=[os.path.join(gh, s) for s in (
                "hierarchy_loader.cpp", "hierarchy_writer.cpp", "traversal.cpp",
                "runtime_switching.cu", "torch/torch_interface.cpp", "ext.cpp")],
            extra_include_paths=[os.path.join(gh, "dependencies/eigen")],
            extra_cflags=["-O3", "-include", "cstdint", "-w"],
            extra_cuda_cflags=["-include", "cstdint", "-w"],
            build_directory=bdir, verbose=verbose, is_python_module=False)
        shutil.copy2(os.path.join(bdir, "ref_gh_C.so"), os.path.join(OUT, "ref_gh_C.so"))
    return ok


if __name__ == "__main__":
    build(tuple(sys.argv[1:]) or ("hr", "gh", "tool"))
    print("[build_ref] done:", sorted(f for f in os.listdir(OUT) if f.endswith(".so") or f == "ref_gh_tool"))

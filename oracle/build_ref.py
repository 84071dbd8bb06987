"""Recipe that compiles the UNMODIFIED reference rasterizer for sm_100a into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  The reference's hot path is CUDA C++ that builds from
five of its own source files; this script runs nvcc on those
files where they lie under /root/reference (nothing is copied into the repo, the
reference's setup.py / CMake are not run) and writes ONE artefact:

    oracle/_ref/_dgr_ref.so      (python extension module, pybind names as in
                                  submodules/diff-gaussian-rasterization/ext.cpp:15-19)

and, for the simple-knn row, a second one from the single source file
submodules/simple-knn/simple_knn.cu (no torch needed; `SimpleKNN::knn` is called through its mangled name):

    oracle/_ref/_sknn_ref.so

oracle/_ref/ is git-ignored but travels to the GPU box with the snapshot.  It cannot
execute in the build container (no GPU); on the GPU box it is
  * the parity pin for keys / sorted order / tile ranges / images / gradients, and
  * the "reference rasterizer" timed by `bench.py --impl reference`.
The only deviation from the reference's own flags is `-include cstdint` (the source
forgets that header, rasterizer_impl.h:24,40) and the explicit sm_100a target.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "_dgr_ref.so")
REF = os.environ.get("GSR_REFERENCE_ROOT", "/root/reference")
DGR = os.path.join(REF, "submodules", "diff-gaussian-rasterization")
SOURCES = ["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
           "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp"]


def available() -> bool:
    return os.path.exists(OUT)


def build(force: bool = False, verbose: bool = False) -> str | None:
    """Returns the path of the built module, or None when /root/reference is absent."""
    if not os.path.isdir(DGR):
        return OUT if available() else None
    if available() and not force:
        newest = max(os.path.getmtime(os.path.join(DGR, s)) for s in SOURCES)
        if os.path.getmtime(OUT) >= newest:
            return OUT
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(OUT_DIR, exist_ok=True)
    build_dir = os.path.join(OUT_DIR, "obj")
    os.makedirs(build_dir, exist_ok=True)
    inc = ["-I" + p for p in ce.include_paths("cuda")] + ["-I" + sysconfig.get_paths()["include"],
           "-I" + os.path.join(DGR, "third_party", "glm"), "-I" + DGR]
    common = ["-std=c++17", "-O3", "-include", "cstdint", "-Xcompiler", "-fPIC",
              "-DTORCH_EXTENSION_NAME=_dgr_ref", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
              "-gencode", "arch=compute_100a,code=sm_100a", "-w"]
    objs = []
    for s in SOURCES:
        o = os.path.join(build_dir, os.path.basename(s) + ".o")
        cmd = ["nvcc", "-c", os.path.join(DGR, s), "-o", o] + common + inc
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
        objs.append(o)
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    link = ["nvcc", "-shared", "-o", OUT] + objs + ["-L" + libdir, "-lc10", "-lc10_cuda", "-ltorch_cpu",
            "-ltorch_cuda", "-ltorch", "-ltorch_python", "-Xlinker", "-rpath", "-Xlinker", libdir]
    if verbose:
        print(" ".join(link))
    subprocess.check_call(link)
    return OUT


SKNN_DIR = os.path.join(REF, "submodules", "simple-knn")
SKNN_OUT = os.path.join(OUT_DIR, "_sknn_ref.so")
SKNN_SYMBOL = "_ZN9SimpleKNN3knnEiP6float3Pf"   # SimpleKNN::knn(int, float3*, float*), simple_knn.h


def build_sknn(force: bool = False, verbose: bool = False) -> str | None:
    """oracle/_ref/_sknn_ref.so from simple_knn.cu alone.  Additions to the reference's flags: `-include cfloat`
    (FLT_MAX is used without its header) and the sm_100a target."""
    src = os.path.join(SKNN_DIR, "simple_knn.cu")
    if not os.path.exists(src):
        return SKNN_OUT if os.path.exists(SKNN_OUT) else None
    if os.path.exists(SKNN_OUT) and not force and os.path.getmtime(SKNN_OUT) >= os.path.getmtime(src):
        return SKNN_OUT
    os.makedirs(OUT_DIR, exist_ok=True)
    cmd = ["nvcc", "-shared", "-o", SKNN_OUT, src, "-std=c++17", "-O3", "-include", "cstdint", "-include", "cfloat",
           "-I" + SKNN_DIR, "-Xcompiler", "-fPIC", "-gencode", "arch=compute_100a,code=sm_100a", "-w"]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return SKNN_OUT


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose=True)
    print("built:" if p else "reference sources not found; nothing built", p or "")
    p = build_sknn(force="--force" in sys.argv, verbose=True)
    print("built:" if p else "simple-knn sources not found; nothing built", p or "")

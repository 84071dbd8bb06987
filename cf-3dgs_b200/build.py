"""Build libgsr_b200.so (the C-ABI library of include/gsr.h) in-tree with nvcc for sm_100a."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libgsr_b200.so")
SOURCES = ["gsr_api.cu", "gsr_preprocess.cu", "gsr_sort.cu", "gsr_render_fwd.cu", "gsr_render_bwd.cu",
           "gsr_preprocess_bwd.cu", "gsr_pose_static.cu", "gsr_loss.cu", "gsr_model.cu", "gsr_knn.cu", "gsr_peaks.cu", "gsr_pose_opt.cu"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]


def _sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "gsr.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, extra=(), out: str = OUT) -> str:
    if not force and not needs_build() and out == OUT:
        return OUT
    objs = []
    obj_dir = os.path.join(HERE, "build", os.path.basename(out))
    os.makedirs(obj_dir, exist_ok=True)
    procs = []
    for src in _sources():
        o = os.path.join(obj_dir, os.path.basename(src) + ".o")
        cmd = ["nvcc", "-c", src, "-o", o] + NVCC_FLAGS + list(extra)
        if verbose:
            print(" ".join(cmd))
        procs.append((cmd, subprocess.Popen(cmd)))
        objs.append(o)
    for cmd, p in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    link = ["nvcc", "-shared", "-o", out] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    if verbose:
        print(" ".join(link))
    subprocess.check_call(link)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True, extra=["-Xptxas", "-v"] if "-v" in sys.argv else ()))

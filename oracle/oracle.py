"""numpy front-end of the C oracle (oracle/gsr_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of gsr_oracle.c.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package (cf-3dgs_b200/) never does.

The call structure mirrors the reference binding
(/root/reference/submodules/diff-gaussian-rasterization/rasterize_points.cu:35-119
forward, :121-208 backward) so parity tests read like calls into the reference.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import POINTER, c_float, c_int, c_int64, c_uint8, c_uint32, c_uint64, c_void_p

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "gsr_oracle.c")
_SO = os.path.join(_HERE, "libgsr_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C oracle (gcc, OpenMP, no FP contraction)."""
    if (not force) and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(_SRC):
        return _SO
    cmd = ["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-mfma",
           "-ffp-contract=off", "-fno-fast-math", _SRC, "-o", _SO, "-lm"]
    subprocess.check_call(cmd)
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.gsro_preprocess.restype = c_int64
        _lib.gsro_bin.restype = c_int
        _lib.gsro_sort_bits.restype = c_int
    return _lib


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _p(a):
    return None if a is None else a.ctypes.data_as(c_void_p)


def sort_bits(W: int, H: int) -> int:
    return int(_load().gsro_sort_bits(c_int(W), c_int(H)))


def forward(means3D, opacities, *, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, viewmatrix, projmatrix, campos, bg, W, H, tanfovx, tanfovy,
            sh_degree=0, scale_modifier=1.0, threads=None, want_unsorted=False):
    """Full forward pass.  Returns a dict holding every intermediate."""
    lib = _load()
    if threads:
        os.environ["OMP_NUM_THREADS"] = str(threads)
    means3D = _f32(means3D); opacities = _f32(opacities).reshape(-1)
    shs = _f32(shs); colors_precomp = _f32(colors_precomp); scales = _f32(scales)
    rotations = _f32(rotations); cov3D_precomp = _f32(cov3D_precomp)
    V = _f32(viewmatrix).reshape(-1); Pm = _f32(projmatrix).reshape(-1)
    campos = _f32(campos).reshape(-1); bg = _f32(bg).reshape(-1)
    P = means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    st = dict(P=P, W=W, H=H, M=M, D=sh_degree)
    st["radii"] = np.zeros(P, np.int32)
    st["xy"] = np.zeros((P, 2), np.float32)
    st["depths"] = np.zeros(P, np.float32)
    st["cov3D"] = np.zeros((P, 6), np.float32)
    st["rgb"] = np.zeros((P, 3), np.float32)
    st["conic_opacity"] = np.zeros((P, 4), np.float32)
    st["tiles_touched"] = np.zeros(P, np.uint32)
    st["clamped"] = np.zeros((P, 3), np.uint8)
    st["rects"] = np.zeros((P, 4), np.uint32)
    R = lib.gsro_preprocess(
        c_int(P), c_int(sh_degree), c_int(M), _p(means3D), _p(scales), c_float(scale_modifier),
        _p(rotations), _p(opacities), _p(shs), _p(cov3D_precomp), _p(colors_precomp), _p(V), _p(Pm),
        _p(campos), c_int(W), c_int(H), c_float(tanfovx), c_float(tanfovy),
        _p(st["radii"]), _p(st["xy"]), _p(st["depths"]), _p(st["cov3D"]), _p(st["rgb"]),
        _p(st["conic_opacity"]), _p(st["tiles_touched"]), _p(st["clamped"]), _p(st["rects"]))
    R = int(R)
    st["num_rendered"] = R
    st["keys_sorted"] = np.zeros(R, np.uint64)
    st["point_list"] = np.zeros(R, np.uint32)
    st["ranges"] = np.zeros((gx * gy, 2), np.uint32)
    ku = np.zeros(R, np.uint64) if want_unsorted else None
    vu = np.zeros(R, np.uint32) if want_unsorted else None
    rc = lib.gsro_bin(c_int(P), c_int(W), c_int(H), _p(st["radii"]), _p(st["rects"]), _p(st["depths"]),
                      _p(st["tiles_touched"]), c_int64(R), _p(st["keys_sorted"]), _p(st["point_list"]),
                      _p(st["ranges"]), _p(ku), _p(vu))
    if rc != 0:
        raise RuntimeError(f"gsro_bin failed: {rc}")
    if want_unsorted:
        st["keys_unsorted"], st["vals_unsorted"] = ku, vu
    feat = colors_precomp if colors_precomp is not None else st["rgb"]
    st["color"] = np.zeros((3, H, W), np.float32)
    st["depth"] = np.zeros((1, H, W), np.float32)
    st["alpha"] = np.zeros((1, H, W), np.float32)
    st["n_contrib"] = np.zeros((H, W), np.uint32)
    lib.gsro_render(c_int(W), c_int(H), _p(st["ranges"]), _p(st["point_list"]), _p(st["xy"]), _p(feat),
                    _p(st["depths"]), _p(st["conic_opacity"]), _p(bg), _p(st["color"]), _p(st["depth"]),
                    _p(st["alpha"]), _p(st["n_contrib"]))
    st["_inputs"] = dict(means3D=means3D, opacities=opacities, shs=shs, colors_precomp=colors_precomp,
                         scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp, V=V, Pm=Pm,
                         campos=campos, bg=bg, tanfovx=tanfovx, tanfovy=tanfovy,
                         scale_modifier=scale_modifier)
    return st


def backward(st, grad_color, grad_depth=None, grad_alpha=None):
    """Backward pass for a state returned by forward().  Returns a dict of gradients
    named like the reference's autograd outputs
    (diff_gaussian_rasterization/__init__.py:146-156)."""
    lib = _load()
    inp = st["_inputs"]
    P, W, H, M, D = st["P"], st["W"], st["H"], st["M"], st["D"]
    grad_color = _f32(grad_color).reshape(3, H, W)
    grad_depth = np.zeros((H, W), np.float32) if grad_depth is None else _f32(grad_depth).reshape(H, W)
    grad_alpha = np.zeros((H, W), np.float32) if grad_alpha is None else _f32(grad_alpha).reshape(H, W)
    feat = inp["colors_precomp"] if inp["colors_precomp"] is not None else st["rgb"]
    g = {}
    g["means2D"] = np.zeros((P, 3), np.float32)
    dconic = np.zeros((P, 4), np.float32)
    g["opacities"] = np.zeros((P, 1), np.float32)
    dcolors = np.zeros((P, 3), np.float32)
    ddepths = np.zeros(P, np.float32)
    lib.gsro_render_backward(
        c_int(P), c_int(W), c_int(H), _p(st["ranges"]), _p(st["point_list"]), _p(st["xy"]), _p(feat),
        _p(st["depths"]), _p(st["conic_opacity"]), _p(inp["bg"]), _p(st["alpha"]), _p(st["n_contrib"]),
        _p(grad_color), _p(grad_depth), _p(grad_alpha), _p(g["means2D"]), _p(dconic), _p(g["opacities"]),
        _p(dcolors), _p(ddepths))
    g["_dL_dconic"] = dconic
    g["_dL_ddepths"] = ddepths
    g["means3D"] = np.zeros((P, 3), np.float32)
    dcov3D = np.zeros((P, 6), np.float32)
    dsh = np.zeros((P, M, 3), np.float32) if inp["shs"] is not None else None
    dscale = np.zeros((P, 3), np.float32) if inp["scales"] is not None else None
    drot = np.zeros((P, 4), np.float32) if inp["scales"] is not None else None
    cov3Ds = inp["cov3D_precomp"] if inp["cov3D_precomp"] is not None else st["cov3D"]
    lib.gsro_preprocess_backward(
        c_int(P), c_int(D), c_int(M), _p(inp["means3D"]), _p(st["radii"]), _p(inp["shs"]),
        _p(st["clamped"]), _p(inp["scales"]), _p(inp["rotations"]), c_float(inp["scale_modifier"]),
        _p(cov3Ds), _p(inp["V"]), _p(inp["Pm"]), _p(inp["campos"]), c_int(W), c_int(H),
        c_float(inp["tanfovx"]), c_float(inp["tanfovy"]), _p(g["means2D"]), _p(dconic), _p(dcolors),
        _p(ddepths), _p(g["means3D"]), _p(dcov3D), _p(dsh), _p(dscale), _p(drot))
    g["colors_precomp"] = dcolors
    g["cov3Ds_precomp"] = dcov3D
    g["sh"] = dsh
    g["scales"] = dscale
    g["rotations"] = drot
    return g


def knn_mean_dist2(points):
    """distCUDA2 (simple-knn): mean squared distance to the three nearest other points, brute force O(P^2)."""
    lib = _load()
    points = _f32(points).reshape(-1, 3)
    out = np.zeros(points.shape[0], np.float32)
    lib.gsro_knn_mean_dist2(c_int(points.shape[0]), _p(points), _p(out))
    return out


def mark_visible(means3D, viewmatrix):
    lib = _load()
    means3D = _f32(means3D)
    V = _f32(viewmatrix).reshape(-1)
    out = np.zeros(means3D.shape[0], np.uint8)
    lib.gsro_mark_visible(c_int(means3D.shape[0]), _p(means3D), _p(V), _p(out))
    return out.astype(bool)

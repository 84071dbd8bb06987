"""Caller for the reference CUDA rasterizer built by oracle/build_ref.py (oracle/_ref/_dgr_ref.so).

TEST INFRASTRUCTURE ONLY.  Calls the reference's own pybind entry points
(submodules/diff-gaussian-rasterization/ext.cpp:15-19) with the argument order of its Python wrapper
(diff_gaussian_rasterization/__init__.py:60-80 forward, :109-132 backward) and parses its opaque state
blobs with the layouts of rasterizer_impl.cu:155-193 (every array starts at the next 128-byte boundary of
the absolute address).  Needs a GPU; never read on the product path.
"""
from __future__ import annotations

import importlib
import os
import sys

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_mod = None


def available() -> bool:
    return os.path.exists(os.path.join(_REF_DIR, "_dgr_ref.so")) and torch.cuda.is_available()


def module():
    global _mod
    if _mod is None:
        if _REF_DIR not in sys.path:
            sys.path.insert(0, _REF_DIR)
        _mod = importlib.import_module("_dgr_ref")
    return _mod


def _e():
    return torch.Tensor([])


def forward(means3D, opacities, *, shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None,
            viewmatrix, projmatrix, campos, bg, W, H, tanfovx, tanfovy, sh_degree=0, scale_modifier=1.0,
            prefiltered=False, debug=False):
    C = module()
    args = (bg, means3D, _e() if colors_precomp is None else colors_precomp, opacities,
            _e() if scales is None else scales, _e() if rotations is None else rotations, float(scale_modifier),
            _e() if cov3D_precomp is None else cov3D_precomp, viewmatrix, projmatrix, float(tanfovx), float(tanfovy),
            int(H), int(W), _e() if shs is None else shs, int(sh_degree), campos, bool(prefiltered), bool(debug))
    R, color, depth, alpha, radii, geom, binning, img = C.rasterize_gaussians(*args)
    return dict(num_rendered=R, color=color, depth=depth, alpha=alpha, radii=radii, geom=geom, binning=binning,
                img=img, _args=args)


def backward(st, grad_color, grad_depth, grad_alpha):
    C = module()
    (bg, means3D, colors, opac, scales, rots, mod, cov3D, view, proj, tx, ty, H, W, sh, deg, campos, _pf, dbg) = st["_args"]
    out = C.rasterize_gaussians_backward(bg, means3D, st["radii"], colors, scales, rots, mod, cov3D, view, proj, tx, ty,
                                         grad_color, grad_depth, grad_alpha, sh, deg, campos, st["geom"],
                                         st["num_rendered"], st["binning"], st["img"], st["alpha"], dbg)
    names = ["means2D", "colors_precomp", "opacities", "means3D", "cov3Ds_precomp", "sh", "scales", "rotations"]
    return dict(zip(names, out))


def _carve(blob: torch.Tensor, offset: int, nbytes: int, dtype):
    base = blob.data_ptr()
    start = ((base + offset + 127) & ~127) - base
    view = blob[start:start + nbytes].view(dtype)
    return view, start + nbytes


def parse_state(st, P, W, H):
    """Sorted keys, point list, ranges, n_contrib and per-Gaussian intermediates from the reference blobs."""
    R = st["num_rendered"]
    out = {}
    b = st["binning"]
    off = 0
    out["point_list"], off = _carve(b, off, 4 * R, torch.int32)
    _, off = _carve(b, off, 4 * R, torch.int32)
    out["keys_sorted"], off = _carve(b, off, 8 * R, torch.int64)
    out["keys_unsorted"], off = _carve(b, off, 8 * R, torch.int64)
    im = st["img"]
    off = 0
    N = W * H
    nc, off = _carve(im, off, 4 * N, torch.int32)
    rg, off = _carve(im, off, 8 * N, torch.int32)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["n_contrib"] = nc.view(H, W)
    out["ranges"] = rg.view(-1, 2)[:tiles]
    g = st["geom"]
    off = 0
    out["depths"], off = _carve(g, off, 4 * P, torch.float32)
    cl, off = _carve(g, off, 3 * P, torch.uint8)
    out["clamped"] = cl.view(P, 3)
    _, off = _carve(g, off, 4 * P, torch.int32)
    m2, off = _carve(g, off, 8 * P, torch.float32)
    out["means2D"] = m2.view(P, 2)
    c3, off = _carve(g, off, 24 * P, torch.float32)
    out["cov3D"] = c3.view(P, 6)
    co, off = _carve(g, off, 16 * P, torch.float32)
    out["conic_opacity"] = co.view(P, 4)
    rgb, off = _carve(g, off, 12 * P, torch.float32)
    out["rgb"] = rgb.view(P, 3)
    out["tiles_touched"], off = _carve(g, off, 4 * P, torch.int32)
    return out


# ---- simple-knn (oracle/_ref/_sknn_ref.so) -------------------------------------------------------
_sknn = None


def sknn_available() -> bool:
    return os.path.exists(os.path.join(_REF_DIR, "_sknn_ref.so")) and torch.cuda.is_available()


def dist_cuda2(points: torch.Tensor) -> torch.Tensor:
    """The reference's distCUDA2 (spatial.cu:15-26): SimpleKNN::knn on a contiguous float32 CUDA tensor."""
    global _sknn
    import ctypes
    if _sknn is None:
        _sknn = ctypes.CDLL(os.path.join(_REF_DIR, "_sknn_ref.so"))
    fn = getattr(_sknn, "_ZN9SimpleKNN3knnEiP6float3Pf")
    fn.restype = None
    fn.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    pts = points.contiguous().float()
    out = torch.zeros(pts.shape[0], dtype=torch.float32, device=pts.device)
    torch.cuda.synchronize()
    fn(pts.shape[0], pts.data_ptr(), out.data_ptr())   # launches on the legacy default stream
    torch.cuda.synchronize()
    return out

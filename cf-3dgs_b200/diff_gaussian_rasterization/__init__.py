"""Drop-in replacement for the `diff_gaussian_rasterization` package CF-3DGS imports
(gaussian_renderer/__init__.py:12, scene/gaussian_model_cf.py:28-31, scene/gaussian_model.py:27-30).

Same public names, argument meaning, return tuple and error behaviour as the reference module
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:17-223); the work is done
by hand-written sm_100a kernels behind the C ABI of include/gsr.h (libgsr_b200.so) -- no CPU path, no
PyTorch fallback.  PyTorch is used for device memory, the current stream and autograd plumbing only.

Put `cf-3dgs_b200/` on sys.path (or PYTHONPATH) ahead of any installed reference build and
`from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer` resolves here.
"""
from __future__ import annotations

from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from . import _cabi
from ._cabi import check, lib

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians",
           "cpu_deep_copy_tuple", "rasterizer_state", "set_sync_free", "sync_free_status"]


def cpu_deep_copy_tuple(input_tuple):
    return tuple(x.cpu().clone() if isinstance(x, torch.Tensor) else x for x in input_tuple)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


# ---------------------------------------------------------------------------------------------------
# Sync-free mode (not in the reference API).  The reference -- and the default path here -- reads num_rendered back to
# the host in every forward to size the binning workspace.  With set_sync_free(True) the forward goes through
# gsr_forward_fused: the binning workspace is sized from a running estimate (last count x slack), the count stays on
# the device, nothing synchronises, and the whole training step can be captured in a CUDA graph.  If a call ever needs
# more room than it was given, its images are incomplete; that is detected from the status word the next time the
# wrapper gets a chance to look (never inside a graph capture), the estimate is raised and a warning is issued.
class _SyncFree:
    enabled = False
    slack = 1.5
    capacity = {}      # device index -> instances the binning workspace is sized for
    pending = {}       # device index -> (event, pinned copy of the status) of a look that is in flight
    last_status = {}   # device index -> (num_rendered, overflowed) of the most recent look
    warned = set()     # devices whose sticky overflow flag has been reported by an in-flight look already
    device_status = {}  # device index -> the ONE status buffer of that device: every sync-free forward (eager, captured
    #                     or replayed, any view) max-es its count and ORs its flags into it; only a look clears it


def set_sync_free(enabled: bool, slack: float = 1.5, capacity: Optional[int] = None, device: Optional[int] = None) -> None:
    """Switch the host-synchronisation-free forward on or off.  `capacity` pins the binning capacity (instances)
    for `device` (default: current); otherwise the first call on a device runs the synchronising path once to learn
    the instance count."""
    _SyncFree.enabled = bool(enabled)
    _SyncFree.slack = float(slack)
    if capacity is not None:
        dev = torch.cuda.current_device() if device is None else device
        _SyncFree.capacity[dev] = int(capacity)


def sync_free_status(device: Optional[int] = None):
    """(largest num_rendered, overflowed) over every sync-free forward on `device` since the previous look -- eager
    calls, captured calls and every replay of a captured graph, whichever view they rendered; synchronises and clears
    the device-side record.  Call it after graph replays: `overflowed` means at least one of those renders was
    incomplete (its farthest instances were dropped) and has to be repeated with the capacity this call has raised."""
    dev = torch.cuda.current_device() if device is None else device
    _SyncFree.pending.pop(dev, None)
    st = _SyncFree.device_status.get(dev)
    if st is not None:
        with torch.cuda.device(dev):
            vals = st.cpu()
            st.zero_()
        _SyncFree.warned.discard(dev)
        _absorb_values(dev, int(vals[0]), int(vals[1]))
        _SyncFree.warned.discard(dev)
    return _SyncFree.last_status.get(dev)


def _status_buffer(dev: int, device: torch.device) -> torch.Tensor:
    st = _SyncFree.device_status.get(dev)
    if st is None:
        st = torch.zeros(2, dtype=torch.int64, device=device)
        _SyncFree.device_status[dev] = st
    return st


def _absorb_status(dev: int) -> None:
    ev, pinned = _SyncFree.pending.pop(dev)
    # the overflow flag is sticky on the device until sync_free_status() clears it (renders queued behind this look may
    # have overflowed too and must not be lost): warn once, keep raising the capacity
    _absorb_values(dev, int(pinned[0]), int(pinned[1]), quiet=dev in _SyncFree.warned)


def _absorb_values(dev: int, R: int, flags: int, quiet: bool = False) -> None:
    _SyncFree.last_status[dev] = (R, bool(flags & 1))
    want = int(R * _SyncFree.slack) + 1024
    if flags & 1 and not quiet:
        _SyncFree.warned.add(dev)
        import warnings
        warnings.warn(f"sync-free rasterizer: {R} tile instances exceeded the binning capacity "
                      f"{_SyncFree.capacity.get(dev)}; that render was incomplete, capacity raised to {want}")
    if want > _SyncFree.capacity.get(dev, 0):
        _SyncFree.capacity[dev] = want


_pinned_counts = {}


def _count_word(device: torch.device) -> torch.Tensor:
    """One pinned int64 per device: the landing spot of num_rendered (rasterizer_impl.cu:282)."""
    key = device.index if device.index is not None else torch.cuda.current_device()
    w = _pinned_counts.get(key)
    if w is None:
        w = torch.zeros(2, dtype=torch.int64).pin_memory()
        _pinned_counts[key] = w
    return w


def _ptr(t: Optional[torch.Tensor]):
    """Device pointer or NULL for the reference's "empty CPU tensor" sentinels."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _dense(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.numel() == 0:
        return t
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32")
    return t.contiguous()


def _forward_impl(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3Ds_precomp,
                  viewmatrix, projmatrix, tanfovx, tanfovy, H, W, sh, degree, campos, prefiltered, debug, rigid=None):
    """Counterpart of _C.rasterize_gaussians (rasterize_points.cu:35-119); same return tuple.
    rigid (optional, 12 floats on the device, row-major 3x4 [R | t]): the kernels move means3D by it first."""
    if means3D.ndim != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    dev = means3D.device
    P = means3D.shape[0]
    means3D = _dense(means3D, "means3D"); colors_precomp = _dense(colors_precomp, "colors_precomp")
    opacities = _dense(opacities, "opacities"); scales = _dense(scales, "scales")
    rotations = _dense(rotations, "rotations"); cov3Ds_precomp = _dense(cov3Ds_precomp, "cov3D_precomp")
    sh = _dense(sh, "sh"); bg = _dense(bg, "bg"); viewmatrix = _dense(viewmatrix, "viewmatrix")
    projmatrix = _dense(projmatrix, "projmatrix"); campos = _dense(campos, "campos")
    M = sh.shape[1] if sh.numel() != 0 else 0
    f32 = dict(dtype=torch.float32, device=dev)
    u8 = dict(dtype=torch.uint8, device=dev)
    color = torch.empty((3, H, W), **f32)
    depth = torch.empty((1, H, W), **f32)
    alpha = torch.empty((1, H, W), **f32)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    if P == 0:
        color.zero_(); depth.zero_(); alpha.zero_()
        e = torch.empty(0, **u8)
        return 0, color, depth, alpha, radii, e, e.clone(), e.clone()
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream().cuda_stream
        geom = torch.empty(lib.gsr_geometry_bytes(P), **u8)
        img = torch.empty(lib.gsr_image_bytes(W, H), **u8)
        di = dev.index if dev.index is not None else torch.cuda.current_device()
        if _SyncFree.enabled and not prefiltered and di in _SyncFree.capacity:
            capturing = torch.cuda.is_current_stream_capturing()
            if not capturing and di in _SyncFree.pending and _SyncFree.pending[di][0].query():
                _absorb_status(di)
            cap = _SyncFree.capacity[di]
            binning = torch.empty(lib.gsr_binning_bytes(cap, W, H), **u8)
            status = _status_buffer(di, dev)
            check(lib.gsr_forward_fused_rigid(
                P, int(degree), int(M), _ptr(means3D), _ptr(scales), float(scale_modifier), _ptr(rotations),
                _ptr(opacities), _ptr(sh), _ptr(cov3Ds_precomp), _ptr(colors_precomp), _ptr(viewmatrix),
                _ptr(projmatrix), _ptr(campos), int(W), int(H), float(tanfovx), float(tanfovy), _ptr(bg),
                radii.data_ptr(), geom.data_ptr(), geom.numel(), binning.data_ptr(), binning.numel(), cap,
                img.data_ptr(), color.data_ptr(), depth.data_ptr(), alpha.data_ptr(), status.data_ptr(),
                None if rigid is None else rigid.data_ptr(), stream, int(bool(debug))))
            if not capturing and di not in _SyncFree.pending:
                pinned = torch.empty(2, dtype=torch.int64).pin_memory()
                pinned.copy_(status, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
                _SyncFree.pending[di] = (ev, pinned)
            return cap, color, depth, alpha, radii, geom, binning, img
        word = _count_word(dev)
        check(lib.gsr_forward_prepare_rigid(
            P, int(degree), int(M), _ptr(means3D), _ptr(scales), float(scale_modifier), _ptr(rotations),
            _ptr(opacities), _ptr(sh), _ptr(cov3Ds_precomp), _ptr(colors_precomp), _ptr(viewmatrix),
            _ptr(projmatrix), _ptr(campos), int(W), int(H), float(tanfovx), float(tanfovy), int(bool(prefiltered)),
            radii.data_ptr(), geom.data_ptr(), geom.numel(), word.data_ptr(),
            None if rigid is None else rigid.data_ptr(), stream, int(bool(debug))))
        R = int(word[0])
        if _SyncFree.enabled:
            _SyncFree.capacity[di] = max(_SyncFree.capacity.get(di, 0), int(R * _SyncFree.slack) + 1024)
        binning = torch.empty(lib.gsr_binning_bytes(R, W, H), **u8)
        check(lib.gsr_forward_render(
            P, int(W), int(H), R, _ptr(bg), _ptr(colors_precomp), geom.data_ptr(), binning.data_ptr(),
            binning.numel(), img.data_ptr(), color.data_ptr(), depth.data_ptr(), alpha.data_ptr(), stream,
            int(bool(debug))))
    return R, color, depth, alpha, radii, geom, binning, img


def _backward_impl(bg, means3D, radii, colors_precomp, scales, rotations, scale_modifier, cov3Ds_precomp,
                   viewmatrix, projmatrix, tanfovx, tanfovy, grad_color, grad_depth, grad_alpha, sh, degree,
                   campos, geom, R, binning, img, alpha, debug, want_pose_grad=False, rigid=None, skip=(), out=None,
                   dsh_M=None):
    """Counterpart of _C.rasterize_gaussians_backward (rasterize_points.cu:121-208).  Returns the same eight
    gradients (None where the matching input was absent) plus the optional fused pose gradient.
    skip: names among {"means2D", "means3D", "opacity", "sh", "colors", "scales", "cov3D"} whose gradients the caller
    does not need (autograd's needs_input_grad): they are neither allocated nor written ("scales" covers rotations).
    out: optional dict of preallocated, contiguous float32 tensors the kernel writes straight into instead of fresh
    ones -- keys "means3D" (P,3), "means2D" (P,3), "opacity" (P,1), "sh" (P,dsh_M,3), "scales" (P,3), "rotations" (P,4):
    slices of one flat buffer in the batched-views loop, so the collective needs no packing copy.
    dsh_M: coefficients per Gaussian in the SH gradient (default M; (degree+1)^2 packs away the inactive bands)."""
    dev = means3D.device
    P = means3D.shape[0]
    H, W = grad_color.shape[1], grad_color.shape[2]
    M = sh.shape[1] if sh.numel() != 0 else 0
    f32 = dict(dtype=torch.float32, device=dev)
    out = out or {}
    dsh_M = M if dsh_M is None else int(dsh_M)

    def _new(name, shape):
        t = out.get(name)
        if t is None:
            return torch.empty(shape, **f32)
        if tuple(t.shape) != tuple(shape) or t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
            raise RuntimeError(f"out[{name!r}] must be a contiguous float32 tensor of shape {tuple(shape)} on {dev}")
        return t

    dmeans3D = _new("means3D", (P, 3)) if "means3D" not in skip else None
    dmeans2D = _new("means2D", (P, 3)) if "means2D" not in skip else None
    dopacity = _new("opacity", (P, 1)) if "opacity" not in skip else None
    dcolors = torch.empty((P, 3), **f32) if colors_precomp.numel() and "colors" not in skip else None
    dsh = _new("sh", (P, dsh_M, 3)) if sh.numel() and "sh" not in skip else None
    dscales = _new("scales", (P, 3)) if scales.numel() and "scales" not in skip else None
    drots = _new("rotations", (P, 4)) if scales.numel() and "scales" not in skip else None
    dcov3D = torch.empty((P, 6), **f32) if cov3Ds_precomp.numel() and "cov3D" not in skip else None
    dpose = torch.empty((6,), **f32) if want_pose_grad else None
    if P == 0:
        if dpose is not None:
            dpose.zero_()
        return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drots, dpose
    grad_color = _dense(grad_color, "grad_color")
    grad_depth = None if grad_depth is None else _dense(grad_depth, "grad_depth")
    grad_alpha = None if grad_alpha is None else _dense(grad_alpha, "grad_alpha")
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream().cuda_stream
        scratch = torch.empty(P * 12, **f32)
        pose_scratch = torch.empty(lib.gsr_pose_scratch_bytes(P) // 4, **f32) if want_pose_grad else None
        check(lib.gsr_backward_rigid(
            P, int(degree), int(M), int(W), int(H), int(R), _ptr(bg), _ptr(means3D), radii.data_ptr(),
            _ptr(colors_precomp), _ptr(scales), _ptr(rotations), float(scale_modifier), _ptr(cov3Ds_precomp),
            _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos), float(tanfovx), float(tanfovy), _ptr(sh),
            alpha.data_ptr(), grad_color.data_ptr(), _ptr(grad_depth), _ptr(grad_alpha), geom.data_ptr(),
            binning.data_ptr(), img.data_ptr(), scratch.data_ptr(), _ptr(dmeans2D), _ptr(dcolors),
            _ptr(dopacity), _ptr(dmeans3D), _ptr(dcov3D), _ptr(dsh), dsh_M, _ptr(dscales), _ptr(drots),
            _ptr(dpose), _ptr(pose_scratch), None if rigid is None else rigid.data_ptr(), stream, int(bool(debug))))
    return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drots, dpose


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # snapshot before anything can be corrupted
            try:
                out = _forward_impl(*args)
            except Exception:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise
        else:
            out = _forward_impl(*args)
        num_rendered, color, depth, alpha, radii, geom, binning, img = out
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.set_materialize_grads(False)   # absent depth/alpha cotangents arrive as None, not zeros
        ctx.mark_non_differentiable(radii)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom,
                              binning, img, alpha)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha):
        rs = ctx.raster_settings
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img, alpha = \
            ctx.saved_tensors
        if grad_color is None:
            grad_color = torch.zeros((3, rs.image_height, rs.image_width), dtype=torch.float32,
                                     device=means3D.device)
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha, sh,
                rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, img, alpha, rs.debug)
        # inputs: means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, settings.  Gradients
        # autograd does not ask for (frozen parameters) are neither allocated nor written by the kernel.
        need = ctx.needs_input_grad
        skip = {n for n, k in (("means3D", need[0]), ("means2D", need[1]), ("sh", need[2]), ("colors", need[3]),
                               ("opacity", need[4]), ("scales", need[5] or need[6]), ("cov3D", need[7])) if not k}
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                res = _backward_impl(*args, skip=skip)
            except Exception:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise
        else:
            res = _backward_impl(*args, skip=skip)
        dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drots, _ = res
        return dmeans3D, dmeans2D, dsh, dcolors, dopacity, dscales, drots, dcov3D, None


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of points in front of the near cull plane (view z > 0.2)."""
        with torch.no_grad():
            rs = self.raster_settings
            positions = _dense(positions, "positions")
            P = positions.shape[0]
            visible = torch.zeros((P,), dtype=torch.bool, device=positions.device)
            if P:
                with torch.cuda.device(positions.device):
                    check(lib.gsr_mark_visible(P, positions.data_ptr(), _dense(rs.viewmatrix, "viewmatrix").data_ptr(),
                                               _dense(rs.projmatrix, "projmatrix").data_ptr(), visible.data_ptr(),
                                               torch.cuda.current_stream().cuda_stream))
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (have_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, rs)


def rasterizer_state(P, W, H, num_rendered, geom, binning, img, actual=None):
    """Parity accessor (not in the reference API): the reference-visible contents of the three workspaces --
    sorted 64-bit keys, point list, tile ranges, n_contrib and the per-Gaussian intermediates -- as tensors.
    After a sync-free forward pass the capacity as num_rendered and the real count as `actual`."""
    dev = geom.device
    R = int(num_rendered)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out = dict(
        keys_sorted=torch.zeros(R, dtype=torch.int64, device=dev),
        point_list=torch.zeros(R, dtype=torch.int32, device=dev),
        ranges=torch.zeros((tiles, 2), dtype=torch.int32, device=dev),
        n_contrib=torch.zeros((H, W), dtype=torch.int32, device=dev),
        depths=torch.zeros(P, dtype=torch.float32, device=dev),
        means2D=torch.zeros((P, 2), dtype=torch.float32, device=dev),
        conic_opacity=torch.zeros((P, 4), dtype=torch.float32, device=dev),
        rgb=torch.zeros((P, 3), dtype=torch.float32, device=dev),
        tiles_touched=torch.zeros(P, dtype=torch.int32, device=dev),
        clamped=torch.zeros((P, 3), dtype=torch.uint8, device=dev),
    )
    if P:
        with torch.cuda.device(dev):
            check(lib.gsr_debug_dump(P, W, H, R, geom.data_ptr(), binning.data_ptr(), img.data_ptr(),
                                     _ptr(out["keys_sorted"]), _ptr(out["point_list"]), out["ranges"].data_ptr(),
                                     out["n_contrib"].data_ptr(), out["depths"].data_ptr(), out["means2D"].data_ptr(),
                                     out["conic_opacity"].data_ptr(), out["rgb"].data_ptr(),
                                     out["tiles_touched"].data_ptr(), out["clamped"].data_ptr(),
                                     torch.cuda.current_stream().cuda_stream))
    if actual is not None:
        out["keys_sorted"] = out["keys_sorted"][:int(actual)]
        out["point_list"] = out["point_list"][:int(actual)]
    return out

"""ctypes binding of libgsr_b200.so (declarations: include/gsr.h).

There is no CPU fallback: if the CUDA library is missing or fails to load, importing this module
raises, and every op built on it fails loudly.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_uint32, c_void_p

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("GSR_B200_LIB", os.path.join(_PKG_ROOT, "libgsr_b200.so"))

# name -> (restype, argtypes); the order mirrors include/gsr.h
_P = c_void_p
SIGNATURES = {
    "gsr_version": (c_int, []),
    "gsr_last_error": (c_char_p, []),
    "gsr_geometry_bytes": (c_size_t, [c_int]),
    "gsr_image_bytes": (c_size_t, [c_int, c_int]),
    "gsr_binning_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "gsr_forward_prepare": (c_int, [c_int, c_int, c_int, _P, _P, c_float, _P, _P, _P, _P, _P, _P, _P, _P,
                                    c_int, c_int, c_float, c_float, c_int, _P, _P, c_size_t, _P, _P, c_int]),
    "gsr_forward_render": (c_int, [c_int, c_int, c_int, c_int64, _P, _P, _P, _P, c_size_t, _P, _P, _P, _P, _P, c_int]),
    "gsr_forward_fused": (c_int, [c_int, c_int, c_int, _P, _P, c_float, _P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int,
                                  c_float, c_float, _P, _P, _P, c_size_t, _P, c_size_t, c_int64, _P, _P, _P, _P, _P, _P,
                                  c_int]),
    "gsr_forward_prepare_rigid": (c_int, [c_int, c_int, c_int, _P, _P, c_float, _P, _P, _P, _P, _P, _P, _P, _P,
                                          c_int, c_int, c_float, c_float, c_int, _P, _P, c_size_t, _P, _P, _P, c_int]),
    "gsr_forward_fused_rigid": (c_int, [c_int, c_int, c_int, _P, _P, c_float, _P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int,
                                        c_float, c_float, _P, _P, _P, c_size_t, _P, c_size_t, c_int64, _P, _P, _P, _P, _P,
                                        _P, _P, c_int]),
    "gsr_backward_rigid": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int64, _P, _P, _P, _P, _P, _P, c_float, _P, _P,
                                   _P, _P, c_float, c_float, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P,
                                   c_int, _P, _P, _P, _P, _P, _P, c_int]),
    "gsr_pose_scratch_bytes": (c_size_t, [c_int]),
    "gsr_backward": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int64, _P, _P, _P, _P, _P, _P, c_float, _P, _P,
                             _P, _P, c_float, c_float, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P,
                             _P, _P, _P, _P, _P, c_int]),
    "gsr_pose_adam_step": (c_int, [_P, _P, _P, c_float, c_float, c_float, c_float, _P]),
    "gsr_pose_gradient_static": (c_int, [c_int, _P, _P, _P, _P]),
    "gsr_photo_loss_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "gsr_photo_loss_forward": (c_int, [c_int, c_int, c_int, _P, _P, c_float, c_int, _P, _P, _P]),
    "gsr_photo_loss_backward": (c_int, [c_int, c_int, c_int, _P, _P, c_float, c_int, _P, _P, _P, _P]),
    "gsr_gaussian_activate": (c_int, [c_int, _P, _P, _P, _P, _P, _P, _P]),
    "gsr_gaussian_adam_state_floats": (c_size_t, [c_int, c_int]),
    "gsr_gaussian_adam_step": (c_int, [c_int, c_int, c_int, c_float, c_float, c_float, _P, _P, _P, _P, _P, _P, _P, _P, _P,
                                       _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gsr_gaussian_adam_step_dev": (c_int, [c_int, c_int, _P, c_float, c_float, c_float, _P, _P, _P, _P, _P, _P, _P, _P, _P,
                                           _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gsr_gaussian_adam_step_part": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P, c_float, c_float, c_float,
                                            _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gsr_gaussian_adam_step_shared": (c_int, [c_int] * 8 + [_P, c_float, c_float, c_float] + [_P] * 18),
    "gsr_densify_stats": (c_int, [c_int, _P, _P, _P, _P, _P, _P]),
    "gsr_multicast_probe": (c_int, [_P, _P, c_int, _P]),
    "gsr_mark_visible": (c_int, [c_int, _P, _P, _P, _P, _P]),
    "gsr_debug_dump": (c_int, [c_int, c_int, c_int, c_int64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gsr_profile_enable": (c_int, [c_int]),
    "gsr_profile_count": (c_int, []),
    "gsr_profile_name": (c_char_p, [c_int]),
    "gsr_profile_read": (c_int, [_P, c_int]),
    "gsr_measure_peaks": (c_int, [_P, _P]),
    "gsr_debug_pair_counts": (c_int, [c_int, c_int, c_int, c_int64, _P, _P, _P, _P, _P]),
    "gsr_knn_scratch_bytes": (c_size_t, [c_int]),
    "gsr_knn_mean_dist2": (c_int, [c_int, _P, _P, _P, c_size_t, _P]),
    "gsr_debug_expf_check": (c_int, [c_uint32, c_uint32, _P, _P]),
}


def load(path: str = LIB_PATH) -> ctypes.CDLL:
    if not os.path.exists(path):
        raise ImportError(
            f"{path} not found: build it with `python cf-3dgs_b200/build.py` (nvcc, sm_100a). "
            "This package has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()


class GsrError(RuntimeError):
    pass


def check(status: int) -> None:
    if status != 0:
        msg = lib.gsr_last_error()
        raise GsrError(f"gsr status {status}: {msg.decode() if msg else ''}")

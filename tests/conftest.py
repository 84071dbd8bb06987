import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "cf-3dgs_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

# Order of the GPU suite (a `-x` run stops at the first failure, so what matters most runs first): golden vectors and
# parity on BASELINE.json's configs, then the size-independent properties (5M stress), the pose / rigid / sync-free
# paths the loop numbers run on, the randomised scenes, the API edge cases, the rows next to the path (model step,
# loss, knn); diagnostics and the one timing comparison near the end; test_zz_* (the no-allocation entry point) last.
_ORDER = ["test_oracle", "test_api_host", "test_cabi", "test_dist_gloo", "test_golden", "test_gpu_parity",
          "test_gpu_properties", "test_gpu_pose", "test_gpu_sync_free", "test_gpu_fuzz", "test_gpu_api",
          "test_gpu_model", "test_gpu_parallel", "test_gpu_loss", "test_gpu_knn", "test_gpu_packed", "test_gpu_timing"]


def _rank(item):
    name = os.path.splitext(os.path.basename(str(item.fspath)))[0]
    if name.startswith("test_zz"):
        return len(_ORDER) + 1
    return _ORDER.index(name) if name in _ORDER else len(_ORDER)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    items.sort(key=_rank)      # stable: the order inside a file is kept
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)

"""GPU: the binning sort has alternative code paths selected by environment variables that are read once per process
(GSR_SORT=onesweep: single-kernel passes with decoupled look-back; GSR_RANK=match|ballot: how lanes holding the same digit
are grouped).  Every combination must produce the same sorted keys, point list and tile ranges, bit for bit."""
import os
import subprocess
import sys

import numpy as np
import pytest

from helpers import ROOT

pytestmark = pytest.mark.gpu

RUNNER = r"""
import sys, numpy as np, torch
sys.path.insert(0, sys.argv[1] + "/cf-3dgs_b200"); sys.path.insert(0, sys.argv[1] + "/tests")
import gsr_synth as S
from helpers import cuda_forward
out = {}
for name, (P, W, H, seed, s0) in dict(a=(40000, 333, 190, 71, None), b=(2500, 64, 64, 72, 0.3), c=(150000, 800, 450, 73, None)).items():
    cam = S.make_camera(W, H); sc = S.make_gvol(P, W, H, seed=seed, s0=s0)
    o, _, _ = cuda_forward(sc, cam, (0, 0, 0))
    st = o["state"]
    out.update({f"{name}_keys": st["keys_sorted"], f"{name}_list": st["point_list"], f"{name}_ranges": st["ranges"],
                f"{name}_color": o["color"]})
np.savez(sys.argv[2], **{k: v.detach().cpu().numpy() for k, v in out.items()})
"""


def _run(env_extra, path):
    env = {k: v for k, v in os.environ.items() if k not in ("GSR_SORT", "GSR_RANK")}
    env.update(env_extra)
    subprocess.run([sys.executable, "-c", RUNNER, ROOT, path], check=True, env=env, timeout=300)
    return np.load(path)


def test_every_sort_mode_gives_the_same_lists(tmp_path):
    base = _run({}, str(tmp_path / "default.npz"))
    for i, extra in enumerate([{"GSR_RANK": "match"}, {"GSR_RANK": "ballot"}, {"GSR_SORT": "onesweep"}]):
        other = _run(extra, str(tmp_path / f"m{i}.npz"))
        for k in base.files:
            assert np.array_equal(base[k].view(np.uint8), other[k].view(np.uint8)), (extra, k)

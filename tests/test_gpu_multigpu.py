"""GPU, two or more devices: the batched-global-views loop under torchrun with the optimiser step shared over NVSwitch
multicast (gsr_gaussian_adam_step_shared) against its NCCL form.  Skipped on a one-GPU box -- the shared step's kernels
are then covered on plain addresses by tests/test_gpu_model.py, the host-side logic by tests/test_dist_gloo.py."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(n, extra, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "examples", "cf3dgs_global.py"), "--batched", "--frames", "4",
           "--points-per-frame", "40000", "--iters", "60", "--width", "480", "--height", "270"] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one NVSwitch / NVLink box")
@pytest.mark.parametrize("deg", [0, 3])
def test_shared_step_over_multicast_trains_like_the_nccl_form(deg):
    n = 2
    nccl = _run(n, ["--sh-degree", str(deg), "--multicast", "off"], 29621)
    mc = _run(n, ["--sh-degree", str(deg), "--multicast", "auto"], 29622)
    assert nccl["replicas_bit_identical"] and mc["replicas_bit_identical"]
    if mc["shared_step_multicast"] != "in use":
        pytest.skip("no multicast on this box: " + mc["shared_step_multicast"])
    # same views, same sums up to their order: the loss curves agree to a fraction of a percent after 60 steps
    assert abs(mc["first_loss"] - nccl["first_loss"]) <= 1e-5 * abs(nccl["first_loss"])
    assert abs(mc["last_loss"] - nccl["last_loss"]) <= 2e-2 * abs(nccl["last_loss"])
    assert mc["last_loss"] < 0.5 * mc["first_loss"]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one NVSwitch / NVLink box")
def test_shared_step_follows_the_sh_degree_schedule_like_the_nccl_form():
    """Degree 0 -> 1 -> 2 during the run: the dense coefficient-0 plane is left, the packing of the SH gradient changes
    twice, the NCCL form switches from one all-reduce to reduce-scatter / all-gather of the SH rows."""
    extra = ["--sh-degree", "2", "--sh-up-every", "20"]
    nccl = _run(2, extra + ["--multicast", "off"], 29623)
    mc = _run(2, extra + ["--multicast", "auto"], 29624)
    assert nccl["replicas_bit_identical"] and mc["replicas_bit_identical"]
    if mc["shared_step_multicast"] != "in use":
        pytest.skip("no multicast on this box: " + mc["shared_step_multicast"])
    assert abs(mc["last_loss"] - nccl["last_loss"]) <= 2e-2 * abs(nccl["last_loss"])
    assert mc["last_loss"] < 0.5 * mc["first_loss"]

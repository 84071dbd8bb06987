"""Golden vectors produced by the reference's own CUDA rasterizer on a B200 (tests/golden/make_golden.py).
CPU part: the C oracle reproduces them (this is what pins the oracle).  GPU part: so does the CUDA path."""
import glob
import os

import numpy as np
import pytest
import torch

from helpers import ROOT

FILES = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "gvol_*.npz")))
KNN_FILES = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "knn_*.npz")))


def _load(path):
    z = np.load(path)
    return {k: z[k] for k in z.files}


def _cam(z):
    P, W, H, deg, R = [int(x) for x in z["dims"]]
    return dict(viewmatrix=z["viewmatrix"], projmatrix=z["projmatrix"], campos=z["campos"], W=W, H=H,
                tanfovx=float(z["tanfov"][0]), tanfovy=float(z["tanfov"][1])), P, W, H, deg, R


@pytest.mark.skipif(not FILES, reason="no golden fixtures committed yet")
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_oracle_reproduces_reference_cuda_output(path):
    from oracle import oracle as O
    z = _load(path)
    cam, P, W, H, deg, R = _cam(z)
    st = O.forward(z["means3D"], z["opacities"], shs=z["shs"], scales=z["scales"], rotations=z["rotations"], bg=z["bg"],
                   sh_degree=deg, **cam)
    assert st["num_rendered"] == R
    assert np.array_equal(st["radii"], z["ref_radii"])
    assert np.array_equal(st["keys_sorted"], z["ref_keys_sorted"].view(np.uint64))
    assert np.array_equal(st["point_list"], z["ref_point_list"].view(np.uint32))
    assert np.array_equal(st["ranges"], z["ref_ranges"].view(np.uint32))
    vis = st["radii"] > 0
    assert np.array_equal(st["depths"].view(np.uint32)[vis], z["ref_depths"].view(np.uint32)[vis])
    assert np.array_equal(st["xy"].view(np.uint32)[vis], z["ref_means2D"].view(np.uint32)[vis])
    assert np.array_equal(st["conic_opacity"].view(np.uint32)[vis], z["ref_conic_opacity"].view(np.uint32)[vis])
    assert np.abs(st["rgb"][vis] - z["ref_rgb"][vis]).max() <= 2e-6
    for k in ["color", "depth", "alpha"]:
        d = np.abs(st[k] - z["ref_" + k])
        assert (d > 1e-5).sum() <= 3, (k, float(d.max()))
    assert (st["n_contrib"] != z["ref_n_contrib"].view(np.uint32)).sum() <= 3
    # backward, fed the reference's own alpha / n_contrib
    st["alpha"] = z["ref_alpha"]
    st["n_contrib"] = z["ref_n_contrib"].view(np.uint32)
    g = O.backward(st, z["grad_color"], z["grad_depth"], z["grad_alpha"])
    for key in ["means3D", "opacities", "sh", "scales", "rotations", "means2D"]:
        a, b = g[key].reshape(-1).astype(np.float64), z["ref_grad_" + key].reshape(-1).astype(np.float64)
        assert np.abs(a - b).max() <= 1e-4 * np.abs(b).max(), key


@pytest.mark.gpu
@pytest.mark.skipif(not FILES, reason="no golden fixtures committed yet")
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_cuda_path_reproduces_reference_cuda_output(path):
    import diff_gaussian_rasterization as D
    z = _load(path)
    cam, P, W, H, deg, R = _cam(z)
    dev = torch.device("cuda:0")
    t = lambda k: torch.from_numpy(z[k]).to(dev)
    leaves = {k: t(k).requires_grad_(True) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    rs = D.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                                         bg=t("bg"), scale_modifier=1.0, viewmatrix=t("viewmatrix"),
                                         projmatrix=t("projmatrix"), sh_degree=deg, campos=t("campos"), prefiltered=False,
                                         debug=False)
    color, radii, depth, alpha = D.GaussianRasterizer(rs)(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"],
                                                          scales=leaves["scales"], rotations=leaves["rotations"])
    fn = color.grad_fn
    st = D.rasterizer_state(P, W, H, fn.num_rendered, fn.saved_tensors[7], fn.saved_tensors[8], fn.saved_tensors[9])
    assert fn.num_rendered == R
    assert np.array_equal(radii.cpu().numpy(), z["ref_radii"])
    assert np.array_equal(st["keys_sorted"].cpu().numpy(), z["ref_keys_sorted"])
    assert np.array_equal(st["point_list"].cpu().numpy(), z["ref_point_list"])
    assert np.array_equal(st["ranges"].cpu().numpy(), z["ref_ranges"])
    assert np.array_equal(st["n_contrib"].cpu().numpy(), z["ref_n_contrib"])
    for k, v in [("color", color), ("depth", depth), ("alpha", alpha)]:
        assert np.abs(v.detach().cpu().numpy() - z["ref_" + k]).max() <= 1e-5, k
    loss = (color * t("grad_color")).sum() + (depth * t("grad_depth")).sum() + (alpha * t("grad_alpha")).sum()
    loss.backward()
    got = {"means3D": leaves["means3D"].grad, "opacities": leaves["opacities"].grad, "sh": leaves["shs"].grad,
           "scales": leaves["scales"].grad, "rotations": leaves["rotations"].grad, "means2D": m2d.grad}
    for key, v in got.items():
        a, b = v.cpu().numpy().reshape(-1).astype(np.float64), z["ref_grad_" + key].reshape(-1).astype(np.float64)
        assert np.abs(a - b).max() <= 1e-4 * np.abs(b).max(), key


# ---- distCUDA2 (tests/golden/make_golden_knn.py: the reference's simple-knn build on a B200) ----
@pytest.mark.skipif(not KNN_FILES, reason="no knn fixtures committed")
@pytest.mark.parametrize("path", KNN_FILES, ids=[os.path.basename(f) for f in KNN_FILES])
def test_knn_oracle_reproduces_reference_output(path):
    """Pins oracle.knn_mean_dist2: bit for bit what SimpleKNN::knn produced."""
    from oracle import oracle as O
    z = _load(path)
    got = O.knn_mean_dist2(z["points"])
    assert np.array_equal(got.view(np.uint32), z["ref_mean_dist2"].view(np.uint32))


@pytest.mark.gpu
@pytest.mark.skipif(not KNN_FILES, reason="no knn fixtures committed")
@pytest.mark.parametrize("path", KNN_FILES, ids=[os.path.basename(f) for f in KNN_FILES])
def test_dist_cuda2_reproduces_reference_output(path):
    from simple_knn._C import distCUDA2
    z = _load(path)
    got = distCUDA2(torch.from_numpy(z["points"]).cuda()).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), z["ref_mean_dist2"].view(np.uint32))

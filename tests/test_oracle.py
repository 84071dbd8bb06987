"""CPU: the C oracle against the independent PyTorch restatement and against its own structural invariants.
(The pin against the reference's CUDA output is tests/test_golden.py.)"""
import numpy as np
import pytest
import torch

import gsr_synth as S
from helpers import np_cam, oracle_forward
from oracle import oracle as O
from oracle import torch_oracle as TO


def _scene(P=2500, W=96, H=64, seed=7, view=0, s0=0.02):
    cam = S.make_camera(W, H) if view == 0 else S.second_view(W, H)
    return S.make_gvol(P, W, H, seed=seed, s0=s0), cam


@pytest.mark.parametrize("view", [0, 1])
@pytest.mark.parametrize("deg", [0, 3])
def test_forward_matches_torch_restatement(view, deg):
    sc, cam = _scene(view=view)
    bg = np.array([1.0, 0.5, 0.2], np.float32)
    st = oracle_forward(sc, cam, bg, sh_degree=deg)
    out = TO.forward(sc["means3D"].double(), sc["opacities"].double(), shs=sc["shs"].double(), scales=sc["scales"].double(),
                     rotations=sc["rotations"].double(), bg=torch.tensor(bg), viewmatrix=cam["viewmatrix"],
                     projmatrix=cam["projmatrix"], campos=cam["campos"], W=cam["W"], H=cam["H"], tanfovx=cam["tanfovx"],
                     tanfovy=cam["tanfovy"], sh_degree=deg)
    assert st["num_rendered"] == out["num_rendered"]
    assert (st["radii"] != out["radii"].numpy()).sum() <= 2
    if (st["radii"] == out["radii"].numpy()).all():
        assert np.array_equal(st["point_list"], out["point_list"].numpy().astype(np.uint32))
        assert np.array_equal(st["ranges"], out["ranges"].numpy().astype(np.uint32))
    for k in ["color", "depth", "alpha"]:
        d = np.abs(st[k] - out[k].detach().numpy())
        assert (d > 2e-5).sum() <= 4, (k, d.max())


@pytest.mark.parametrize("view", [0, 1])
def test_backward_matches_float64_autograd(view):
    sc, cam = _scene(view=view)
    bg = np.array([0.3, 0.6, 0.9], np.float32)
    st = oracle_forward(sc, cam, bg, sh_degree=3)
    leaves = {k: sc[k].double().clone().requires_grad_(True) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    out = TO.forward(leaves["means3D"], leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"],
                     rotations=leaves["rotations"], bg=torch.tensor(bg), viewmatrix=cam["viewmatrix"],
                     projmatrix=cam["projmatrix"], campos=cam["campos"], W=cam["W"], H=cam["H"], tanfovx=cam["tanfovx"],
                     tanfovy=cam["tanfovy"], sh_degree=3)
    ct = S.cotangents(cam["W"], cam["H"], seed=3)
    loss = (out["color"] * ct["grad_color"].double()).sum() + (out["depth"] * ct["grad_depth"].double()).sum() + \
        (out["alpha"] * ct["grad_alpha"].double()).sum()
    loss.backward()
    g = O.backward(st, ct["grad_color"].numpy(), ct["grad_depth"].numpy(), ct["grad_alpha"].numpy())
    for name, key in [("means3D", "means3D"), ("opacities", "opacities"), ("shs", "sh"), ("scales", "scales"),
                      ("rotations", "rotations")]:
        a = leaves[name].grad.numpy().reshape(-1)
        b = g[key].reshape(-1).astype(np.float64)
        # float32 recurrences (T/(1-alpha)) against exact float64: 1e-3 of the largest entry
        assert np.abs(a - b).max() <= 1e-3 * np.abs(a).max(), name


def test_structural_invariants():
    sc, cam = _scene(P=6000, W=130, H=70, seed=5)   # ragged: 130x70 is not a multiple of 16
    st = oracle_forward(sc, cam, np.zeros(3, np.float32))
    R = st["num_rendered"]
    assert R == int(st["tiles_touched"].sum())
    assert (np.diff(st["keys_sorted"].astype(np.uint64)) >= 0).all() if R > 1 else True
    tiles = st["keys_sorted"] >> np.uint64(32)
    rg = st["ranges"]
    for t in np.unique(tiles.astype(np.int64))[:50]:
        a, b = rg[t]
        assert (tiles[a:b] == t).all() and (a == 0 or tiles[a - 1] != t) and (b == R or tiles[b] != t)
    empty = np.setdiff1d(np.arange(rg.shape[0]), np.unique(tiles.astype(np.int64)))
    assert (rg[empty] == 0).all()
    vis = st["radii"] > 0
    assert (st["tiles_touched"][~vis] == 0).all() and (st["tiles_touched"][vis] > 0).all()
    assert st["alpha"].min() >= 0 and st["alpha"].max() <= 1.0 + 1e-5
    assert O.sort_bits(640, 360) == 42 and O.sort_bits(960, 540) == 43 and O.sort_bits(1920, 1080) == 45


def test_all_culled_and_empty():
    sc, cam = _scene(P=300)
    sc["means3D"][:, 2] = -1.0   # everything behind the camera
    bg = np.array([0.1, 0.2, 0.3], np.float32)
    st = oracle_forward(sc, cam, bg)
    assert st["num_rendered"] == 0 and (st["radii"] == 0).all()
    assert np.allclose(st["color"], bg[:, None, None]) and (st["alpha"] == 0).all()
    g = O.backward(st, np.ones((3, cam["H"], cam["W"]), np.float32))
    assert all((np.asarray(v) == 0).all() for k, v in g.items() if v is not None)


def test_mark_visible_is_the_near_plane_test():
    sc, cam = _scene(P=1000)
    vis = O.mark_visible(sc["means3D"].numpy(), cam["viewmatrix"].numpy())
    z = (torch.cat([sc["means3D"], torch.ones(1000, 1)], 1) @ cam["viewmatrix"])[:, 2].numpy()
    assert (vis == (z > 0.2)).mean() > 0.999


def test_knn_oracle_is_the_three_nearest_neighbour_mean():
    """gsro_knn_mean_dist2 (restating simple_knn.cu:131-183) against scipy's exact k-d tree in float64."""
    from scipy.spatial import cKDTree
    from oracle import oracle as O
    rng = np.random.default_rng(5)
    pts = (rng.standard_normal((3000, 3)) * 2 + 1).astype(np.float32)
    got = O.knn_mean_dist2(pts)
    d, _ = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=4)
    want = (d[:, 1:] ** 2).mean(1)
    assert np.abs(got - want).max() <= 1e-6 * want.max()
    # fewer than three neighbours: the missing ones stay at FLT_MAX (simple_knn.cu:151,182)
    assert np.allclose(O.knn_mean_dist2(pts[:3]), np.float32(3.402823466e38) / 3, rtol=1e-6)
    assert np.isinf(O.knn_mean_dist2(pts[:2])).all()


def test_simple_knn_shim_has_no_cpu_path():
    import torch
    from simple_knn._C import distCUDA2
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(8, 3))

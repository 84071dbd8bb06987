"""GPU: distCUDA2 (simple-knn drop-in, csrc/gsr_knn.cu through the C ABI) against the brute-force CPU oracle and, when it
travelled to the box, the reference's own CUDA build (oracle/_ref/_sknn_ref.so).  The output is a selection problem
evaluated with a fixed rounding sequence, so the bar is bit-exact."""
import numpy as np
import pytest
import torch

import gsr_synth as S

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _clouds():
    rng = np.random.default_rng(7)
    yield "tiny1", rng.standard_normal((1, 3))
    yield "tiny2", rng.standard_normal((2, 3))
    yield "tiny3", rng.standard_normal((3, 3))
    yield "tiny4", rng.standard_normal((4, 3))
    yield "tiny5", rng.standard_normal((5, 3))
    yield "normal_1000", rng.standard_normal((1000, 3))
    yield "box_edge_1025", rng.random((1025, 3)) * 4 - 1          # one point over a box boundary
    yield "uniform_5000", rng.random((5000, 3)) * 10 + 3           # away from the origin (the bounding box contains it)
    g = np.stack(np.meshgrid(np.arange(12.0), np.arange(12.0), np.arange(12.0)), -1).reshape(-1, 3)
    yield "lattice_ties", g * 0.25                                   # every distance tied many times over
    dup = rng.standard_normal((700, 3))
    yield "duplicates", np.concatenate([dup, dup[:350], dup[:100]])  # coincident points: zero distances count
    flat = rng.standard_normal((3000, 3)); flat[:, 2] = 0.0
    yield "planar", flat                                             # degenerate axis
    cl = np.concatenate([rng.standard_normal((2000, 3)) * 0.01 + c for c in rng.standard_normal((4, 3)) * 5])
    yield "clusters", cl
    yield "negative_octant", -np.abs(rng.standard_normal((2500, 3))) - 2.0


@pytest.mark.parametrize("name,pts", list(_clouds()), ids=[n for n, _ in _clouds()])
def test_dist_cuda2_matches_the_brute_force_oracle_bit_for_bit(name, pts):
    from oracle import oracle as O
    from simple_knn._C import distCUDA2
    pts = np.ascontiguousarray(pts, np.float32)
    want = O.knn_mean_dist2(pts)
    got = distCUDA2(torch.from_numpy(pts).to(DEV)).cpu().numpy()
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (name, np.abs(got - want).max())


def test_dist_cuda2_matches_the_reference_build_at_model_sizes():
    from oracle import ref_cuda as R
    if not R.sknn_available():
        pytest.skip("oracle/_ref/_sknn_ref.so did not travel to this box")
    from simple_knn._C import distCUDA2
    rng = np.random.default_rng(11)
    clouds = {
        "gvol_300k": S.make_gvol(300_000, 960, 540, seed=3)["means3D"].numpy(),
        "normal_200k": rng.standard_normal((200_000, 3)).astype(np.float32),
        "depth_map_like": np.stack([(np.arange(518_400) % 960) / 960.0, (np.arange(518_400) // 960) / 540.0,
                                    2.0 + 0.3 * np.sin(np.arange(518_400) * 1e-3)], 1).astype(np.float32),
    }
    for name, pts in clouds.items():
        t = torch.from_numpy(np.ascontiguousarray(pts, np.float32)).to(DEV)
        want = R.dist_cuda2(t)
        got = distCUDA2(t)
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)), (name, float((got - want).abs().max()))


def test_dist_cuda2_is_stream_ordered_and_repeatable():
    from simple_knn._C import distCUDA2
    pts = torch.randn(50_000, 3, device=DEV)
    a = distCUDA2(pts)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        b = distCUDA2(pts)
    s.synchronize()
    assert torch.equal(a, b)
    assert torch.equal(distCUDA2(torch.zeros(0, 3, device=DEV)), torch.zeros(0, device=DEV))


def test_dist_cuda2_argument_checks():
    from simple_knn._C import distCUDA2
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(10, 3))                                  # CPU tensor
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(10, 3, device=DEV, dtype=torch.float64))
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(10, 2, device=DEV))

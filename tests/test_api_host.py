"""CPU: host-side behaviour of the drop-in module mirrors the reference wrapper
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:160-223)."""
import pytest
import torch

import diff_gaussian_rasterization as D


def test_public_names():
    for n in ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians",
              "cpu_deep_copy_tuple"]:
        assert hasattr(D, n)


def test_settings_fields_match_reference():
    assert D.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def _rast():
    e = torch.eye(4)
    rs = D.GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, e, e, 0, torch.zeros(3), False, False)
    return D.GaussianRasterizer(rs)


def test_exactly_one_colour_source():
    r = _rast()
    x = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(x, x, torch.zeros(4, 1), scales=x, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(x, x, torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), colors_precomp=x, scales=x, rotations=torch.zeros(4, 4))


def test_exactly_one_covariance_source():
    r = _rast()
    x = torch.zeros(4, 3)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(x, x, torch.zeros(4, 1), colors_precomp=x)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(x, x, torch.zeros(4, 1), colors_precomp=x, scales=x)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(x, x, torch.zeros(4, 1), colors_precomp=x, scales=x, rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))


def test_bad_means_shape_is_a_runtime_error():
    r = _rast()
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        r(torch.zeros(4, 2), torch.zeros(4, 2), torch.zeros(4, 1), colors_precomp=torch.zeros(4, 3),
          scales=torch.zeros(4, 3), rotations=torch.zeros(4, 4))


def test_cpu_inputs_are_rejected_not_silently_computed():
    r = _rast()
    x = torch.rand(4, 3)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        r(x, x, torch.zeros(4, 1), colors_precomp=x, scales=x, rotations=torch.zeros(4, 4))


def test_cpu_deep_copy_tuple():
    t = (torch.ones(2), 3, "s")
    c = D.cpu_deep_copy_tuple(t)
    assert c[1] == 3 and c[2] == "s" and torch.equal(c[0], t[0]) and c[0] is not t[0]


def test_the_product_package_never_touches_the_oracle():
    """oracle/ is the checker: nothing under cf-3dgs_b200/ (the shipped package and its CUDA sources) may import, load or
    name it.  (tools/ and the `--impl reference` arms of bench.py / examples/ time the reference build beside the product;
    they are measurement harnesses, not the product path.)"""
    import os
    import re
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "cf-3dgs_b200")
    # imports, dlopen targets and #includes (comments may cite the oracle by name)
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|import_module\(.*oracle|CDLL\(.*(oracle|_ref)|#\s*include\s*[\"<].*oracle|"
                     r"libgsr_oracle|_dgr_ref|_sknn_ref", re.M)
    offenders = []
    for d, _, files in os.walk(root):
        if os.sep + "build" in d or os.sep + "variants" in d:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                if pat.search(open(os.path.join(d, f), errors="ignore").read()):
                    offenders.append(os.path.join(d, f))
    assert not offenders, offenders


def test_coefficient_0_plane_bookkeeping_of_the_fused_model():
    """FusedGaussianModel.shs() / .features (host logic only, CPU tensors): at degree 0 a dense (P,1,3) plane is handed
    out and is the truth for coefficient 0; reading `features` brings it in; a higher degree or assigning `features` ends
    the mode.  (The kernels that update the plane are covered by tests/test_gpu_model.py.)"""
    import torch
    import gsr_model
    P, M = 7, 16
    g = torch.Generator().manual_seed(0)
    feats = torch.randn(P, M, 3, generator=g)
    lrs = dict(xyz=1e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)
    m = gsr_model.FusedGaussianModel(torch.randn(P, 3), feats, torch.randn(P, 1), torch.randn(P, 3), torch.randn(P, 4), lrs)
    assert not m.dc_mode and torch.equal(m.features, feats)
    sh0 = m.shs(0)
    assert m.dc_mode and sh0.shape == (P, 1, 3) and sh0.is_contiguous() and torch.equal(sh0[:, 0], feats[:, 0])
    assert m.shs(0) is sh0                              # stays the same tensor while the mode lasts
    sh0.add_(1.0)                                       # what a step at degree 0 does: it updates the plane
    full = m.features                                   # reading brings coefficient 0 up to date, the rest is untouched
    assert torch.equal(full[:, 0], feats[:, 0] + 1.0) and torch.equal(full[:, 1:], feats[:, 1:]) and m.dc_mode
    sh2 = m.shs(2)                                      # a higher degree renders from the full tensor and ends the mode
    assert not m.dc_mode and sh2.shape == (P, M, 3) and torch.equal(sh2[:, 0], feats[:, 0] + 1.0)
    m.shs(0).mul_(2.0)
    m.features = torch.zeros(P, M, 3)                   # assigning a new tensor makes that tensor the truth
    assert not m.dc_mode and float(m.features.abs().sum()) == 0.0
    one = gsr_model.FusedGaussianModel(torch.randn(P, 3), feats[:, :1], torch.randn(P, 1), torch.randn(P, 3),
                                       torch.randn(P, 4), lrs)
    assert one.shs(0) is one.features and not one.dc_mode          # a one-coefficient model is dense already

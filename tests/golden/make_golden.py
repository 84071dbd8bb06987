"""Generates the golden vectors under tests/golden/ by running the REFERENCE's own CUDA rasterizer
(oracle/_ref/_dgr_ref.so, built by oracle/build_ref.py from the unmodified sources under /root/reference)
on a B200.  The reference ships no fixtures of its own, so these are the pin:

    gpurun -- python tests/golden/make_golden.py gpurun_out/golden     # on the GPU box
    cp gpurun_out/golden/*.npz tests/golden/                           # back in the repo

Each file holds the seeded inputs and everything the reference produced for them: radii, num_rendered, the sorted
64-bit keys, point list, tile ranges, n_contrib, the three images and the eight gradients.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import gsr_synth as S  # noqa: E402
from oracle import ref_cuda as RC  # noqa: E402

CASES = {
    "gvol_a": dict(P=1500, W=64, H=48, seed=21, view=0, bg=(0.0, 0.0, 0.0), deg=3, s0=0.03, dense=True),
    "gvol_b": dict(P=1500, W=80, H=56, seed=22, view=1, bg=(1.0, 1.0, 1.0), deg=2, s0=0.05, dense=True),
    "gvol_c": dict(P=1000, W=70, H=50, seed=23, view=0, bg=(0.2, 0.5, 0.7), deg=0, s0=0.02, dense=False),
}


def make_case(c):
    cam = S.make_camera(c["W"], c["H"]) if c["view"] == 0 else S.second_view(c["W"], c["H"])
    sc = S.make_gvol(c["P"], c["W"], c["H"], seed=c["seed"], s0=c["s0"], sh_degree=c["deg"])
    ct = S.cotangents(c["W"], c["H"], seed=c["seed"] + 1000, dense=c["dense"])
    return cam, sc, ct


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    dev = torch.device("cuda:0")
    for name, c in CASES.items():
        cam, sc, ct = make_case(c)
        bg = torch.tensor(c["bg"])
        st = RC.forward(sc["means3D"].to(dev), sc["opacities"].to(dev), shs=sc["shs"].to(dev), scales=sc["scales"].to(dev),
                        rotations=sc["rotations"].to(dev), viewmatrix=cam["viewmatrix"].to(dev),
                        projmatrix=cam["projmatrix"].to(dev), campos=cam["campos"].to(dev), bg=bg.to(dev), W=c["W"], H=c["H"],
                        tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=c["deg"])
        torch.cuda.synchronize()
        ps = RC.parse_state(st, c["P"], c["W"], c["H"])
        g = RC.backward(st, ct["grad_color"].to(dev), ct["grad_depth"].to(dev), ct["grad_alpha"].to(dev))
        torch.cuda.synchronize()
        out = dict(
            means3D=sc["means3D"], scales=sc["scales"], rotations=sc["rotations"], opacities=sc["opacities"], shs=sc["shs"],
            viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"], campos=cam["campos"], bg=bg,
            tanfov=torch.tensor([cam["tanfovx"], cam["tanfovy"]], dtype=torch.float64),
            dims=torch.tensor([c["P"], c["W"], c["H"], c["deg"], st["num_rendered"]]),
            grad_color=ct["grad_color"], grad_depth=ct["grad_depth"], grad_alpha=ct["grad_alpha"],
            ref_radii=st["radii"], ref_color=st["color"], ref_depth=st["depth"], ref_alpha=st["alpha"],
            ref_keys_sorted=ps["keys_sorted"], ref_point_list=ps["point_list"], ref_ranges=ps["ranges"],
            ref_n_contrib=ps["n_contrib"], ref_depths=ps["depths"], ref_means2D=ps["means2D"],
            ref_conic_opacity=ps["conic_opacity"], ref_rgb=ps["rgb"], ref_tiles_touched=ps["tiles_touched"],
            ref_clamped=ps["clamped"],
            **{"ref_grad_" + k: v for k, v in g.items()})
        np.savez_compressed(os.path.join(out_dir, name + ".npz"),
                            **{k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in out.items()})
        print(name, "R =", st["num_rendered"], "visible =", int((st["radii"] > 0).sum()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))

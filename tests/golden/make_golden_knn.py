"""Golden vectors for distCUDA2: the REFERENCE's own simple-knn build (oracle/_ref/_sknn_ref.so, compiled by
oracle/build_ref.py from submodules/simple-knn/simple_knn.cu) run on a B200.

    gpurun -- python tests/golden/make_golden_knn.py gpurun_out/golden
    cp gpurun_out/golden/knn_*.npz tests/golden/

Each file: the seeded points and the reference's mean squared 3-NN distance for them."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import gsr_synth as S  # noqa: E402
from oracle import ref_cuda as RC  # noqa: E402


def clouds():
    rng = np.random.default_rng(101)
    yield "knn_normal", rng.standard_normal((6000, 3)).astype(np.float32)
    dup = rng.standard_normal((2500, 3)).astype(np.float32)
    yield "knn_duplicates_offset", (np.concatenate([dup, dup[:1200], dup[:300]]) + np.float32(7.5)).astype(np.float32)
    yield "knn_gpix", S.make_gpix(160, 90, seed=4, target_P=8000)["means3D"].numpy().astype(np.float32)


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    for name, pts in clouds():
        ref = RC.dist_cuda2(torch.from_numpy(pts).cuda()).cpu().numpy()
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), points=pts, ref_mean_dist2=ref)
        print(name, pts.shape, float(ref.mean()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))

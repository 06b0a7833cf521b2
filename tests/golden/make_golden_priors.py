"""Golden fixture for the depth-prior normalisation from the REAL reference (build container only):
    PYTHONPATH=/root/reference python tests/golden/make_golden_priors.py
runs `normalize_depth` (src/models/utils/priors.py:107) on oracle/priors_oracle.py's seeded cases."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import priors_oracle as PO  # noqa: E402


def main():
    from src.models.utils.priors import normalize_depth  # reference

    x = PO.synthetic_depth_cases(seed=0)
    with torch.no_grad():
        y = normalize_depth(torch.from_numpy(x.copy())).numpy()
    np.savez_compressed(os.path.join(HERE, "priors_depth.npz"), out=y, seed=0, torch=torch.__version__)
    print("wrote priors_depth.npz", y.shape, float(y.mean()))

    # pose / ray priors exactly as WorldMirror.extract_priors computes them (worldmirror.py:228-249)
    from src.models.utils.camera_utils import extrinsics_to_vector  # reference
    from src.models.utils.priors import normalize_poses  # reference

    fx = {}
    for i, (pose, intr) in enumerate(PO.synthetic_pose_cases(seed=0)):
        views = {"camera_pose": torch.from_numpy(pose.copy())[None], "camera_intrinsics": torch.from_numpy(intr.copy())[None]}
        h = w = 518
        with torch.no_grad():
            extrinsics = views["camera_pose"][:, :, :3]
            extrinsics = normalize_poses(extrinsics)
            poses = extrinsics_to_vector(extrinsics)[:, :, :7]
            k = views["camera_intrinsics"][:, :, :3, :3]
            rays = torch.stack([k[:, :, 0, 0] / w, k[:, :, 1, 1] / h, k[:, :, 0, 2] / w, k[:, :, 1, 2] / h], dim=-1)
        fx[f"poses{i}"], fx[f"rays{i}"] = poses[0].numpy(), rays[0].numpy()
    np.savez_compressed(os.path.join(HERE, "priors_pose.npz"), **fx)
    print("wrote priors_pose.npz", {k: v.shape for k, v in fx.items()})


if __name__ == "__main__":
    main()

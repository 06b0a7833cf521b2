"""Generate the rasterisation fixture from gsplat's own pure-torch statement of projection and tile binning (run in the
build container only; gsplat's CUDA extension cannot be built here, its torch functions run on the CPU).

    PYTHONPATH=/root/reference:/root/reference/submodules/gsplat python tests/golden/make_golden_raster.py

Stores, for the seeded scene of oracle/raster_oracle.py `synthetic_scene`: covariances, projected means / conics /
depths / radii (fixed 3.33-sigma extent: the torch statement has no opacity-aware extent), tiles per gaussian, sorted
intersection ids, flatten ids and tile offsets.  Pins oracle/raster_oracle.py `project(opacity_aware=False)` and `bin_tiles`.
"""
import math
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import raster_oracle as R  # noqa: E402  (only for the seeded inputs)


def main():
    from gsplat.cuda._torch_impl import (_fully_fused_projection, _isect_offset_encode, _isect_tiles,  # reference dependency
                                         _quat_scale_to_covar_preci)

    sc = R.synthetic_scene()
    W, H = sc["W"], sc["H"]
    t = {k: torch.from_numpy(v) for k, v in sc.items() if isinstance(v, np.ndarray)}
    viewmats = torch.linalg.inv(t["camtoworlds"])
    covars, _ = _quat_scale_to_covar_preci(t["quats"], t["scales"], True, False, triu=False)
    radii, means2d, depths, conics, _ = _fully_fused_projection(t["means"], covars, viewmats, t["Ks"], W, H, eps2d=0.3,
                                                                near_plane=0.01, far_plane=1e10)
    tw, th = math.ceil(W / 16.0), math.ceil(H / 16.0)
    per, ids, flat = _isect_tiles(means2d, radii, depths, 16, tw, th)
    offs = _isect_offset_encode(ids, viewmats.shape[0], tw, th)
    fx = {"covars": covars.numpy(), "radii": radii.numpy(), "means2d": means2d.numpy(), "depths": depths.numpy(),
          "conics": conics.numpy(), "tiles_per_gauss": per.numpy(), "isect_ids": ids.numpy(), "flatten_ids": flat.numpy(),
          "offsets": offs.numpy(), "viewmats": viewmats.numpy(), "torch": torch.__version__}
    print("visible (camera, gaussian) pairs", int((radii > 0).all(-1).sum()), "of", radii.shape[0] * radii.shape[1], "intersections", len(ids))
    path = os.path.join(HERE, "raster_scene.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB")


if __name__ == "__main__":
    main()

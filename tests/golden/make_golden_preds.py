"""Fixture of the WHOLE reference model's forward (run in the build container only).

    PYTHONPATH=/root/reference:/root/reference/submodules/gsplat python tests/golden/make_golden_preds.py

Builds the real `WorldMirror` (src/models/models/worldmirror.py:16) with the two-kwarg constructor shim of SURVEY.md
section 9-D1 (`enable_interpolation`, `max_resolution` are passed to a class that does not take them) and
`enable_gs=False` (gsplat's CUDA extension cannot be built here), loads the synthetic aggregator + head weights with
strict=True and runs `WorldMirror.forward(views, cond_flags)` on config C1 (2 views 252x252, no priors) and on the same
views with all three priors.  Stored: the exact key set of the returned dictionary, `camera_params`, `camera_poses`
(camera-to-world, worldmirror.py:161-175) and `camera_intrs`, and the dense predictions on a stride-6 lattice.  This
pins the public API of `B200WorldMirror` (key set, shapes, the camera matrices) against the reference's own forward.
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
W = importlib.import_module("hunyuanworld-mirror_b200.weights")


def main():
    import src.models.models.worldmirror as wm  # reference

    V = wm.VisualGeometryTransformer
    wm.VisualGeometryTransformer = lambda *a, enable_interpolation=False, max_resolution=2044, **k: V(*a, **k)
    torch.manual_seed(0)
    model = wm.WorldMirror(enable_gs=False).eval()
    sd = {"visual_geometry_transformer." + k: v for k, v in W.synthetic_state_dict(seed=0).items()}
    sd.update(W.synthetic_heads_state_dict(seed=0))
    print(model.load_state_dict(sd, strict=True))
    n, h, w = 2, 252, 252
    imgs = W.synthetic_images(n, h, w, seed=1)
    pose, depth, intr = W.synthetic_scene(n, h, w, seed=2)
    fx = {"n_views": n, "H": h, "W": w, "stride": 6, "torch": torch.__version__}
    for tag, views, flags in (("nopriors", {"img": imgs}, [0, 0, 0]),
                              ("priors", {"img": imgs, "camera_pose": pose, "depthmap": depth, "camera_intrinsics": intr}, [1, 1, 1])):
        with torch.no_grad():
            preds = model(views, cond_flags=flags)
        fx[tag + "_keys"] = np.array(sorted(preds.keys()))
        fx[tag + "_shapes"] = np.array([str(tuple(preds[k].shape)) for k in sorted(preds.keys())])
        for k in ("camera_params", "camera_poses", "camera_intrs"):
            fx[f"{tag}_{k}"] = preds[k].numpy()
        for k in ("depth", "depth_conf", "pts3d", "pts3d_conf", "normals", "normals_conf"):
            fx[f"{tag}_{k}_sub"] = preds[k][0, :, ::6, ::6].numpy()
        print(tag, list(fx[tag + "_keys"]), "fov", preds["camera_params"][0, :, 7:].tolist())
    path = os.path.join(HERE, "c1_preds.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB")


if __name__ == "__main__":
    main()

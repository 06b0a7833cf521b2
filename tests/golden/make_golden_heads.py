"""Generate the prediction-head fixture from the REAL reference (run in the build container only).

    PYTHONPATH=/root/reference:/root/reference/submodules/gsplat python tests/golden/make_golden_heads.py [c1|c2]

Config C1 (2 views 252x252, no priors): the reference `VisualGeometryTransformer` (synthetic aggregator weights)
produces `token_list`; the reference `CameraHead` and three `DPTHead`s (synthetic head weights, strict=True),
constructed exactly as `WorldMirror._init_heads` does (worldmirror.py:66-98), turn it into camera_params,
depth, pts3d and normals in fp32 as `_gen_all_preds` does (worldmirror.py:154-196).  `WorldMirror()` itself
cannot be constructed as shipped, so the modules are built directly.
Stored: camera_params of all four refinement steps, the camera tokens, and the dense predictions on a
stride-3 pixel lattice plus whole-map statistics.  This fixture pins oracle/heads_oracle.py and is the
end-to-end target of the GPU parity test for the tolerances BASELINE.json states.
`c2` = BASELINE.json configs[1], the configuration bench.py times (8 views 518x518, no priors): same recipe on a
stride-7 lattice -> c2_heads.npz.
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
STRIDE = 3


def main():
    from src.models.heads.camera_head import CameraHead  # reference
    from src.models.heads.dense_head import DPTHead  # reference
    from src.models.models.visual_transformer import VisualGeometryTransformer  # reference

    case = sys.argv[1] if len(sys.argv) > 1 else "c1"
    n, h, w, stride = {"c1": (2, 252, 252, STRIDE), "c2": (8, 518, 518, 7)}[case]
    agg = VisualGeometryTransformer(enable_cond=True).eval()
    agg.load_state_dict(W.synthetic_state_dict(seed=0), strict=True)
    hsd = W.synthetic_heads_state_dict(seed=0)

    def sub(p):
        return {k[len(p):]: v for k, v in hsd.items() if k.startswith(p)}

    cam = CameraHead(dim_in=2048).eval()
    cam.load_state_dict(sub("cam_head."), strict=True)
    heads = {}
    for name, pre, od, act in (("pts3d", "pts_head.", 4, "inv_log+expp1"), ("depth", "depth_head.", 2, "exp+expp1"),
                               ("normals", "norm_head.", 4, "norm+expp1")):
        m = DPTHead(dim_in=2048, output_dim=od, patch_size=14, activation=act).eval()
        m.load_state_dict(sub(pre), strict=True)
        heads[name] = m

    imgs = W.synthetic_images(n, h, w, seed=1)
    fx = {"n_views": n, "H": h, "W": w, "stride": stride, "torch": torch.__version__}
    with torch.no_grad():
        token_list, psi = agg(imgs)
        seq = cam(token_list)
        fx["camera_params_steps"] = torch.stack(seq, 0).numpy()  # [4, 1, S, 9]
        fx["cam_tokens"] = token_list[-1][:, :, 0].numpy()
        for name, m in heads.items():
            pred, conf = m([t.clone() for t in token_list], images=imgs, patch_start_idx=psi)
            fx[name + "_sub"] = pred[0, :, ::stride, ::stride].numpy()
            fx[name + "_conf_sub"] = conf[0, :, ::stride, ::stride].numpy()
            fx[name + "_absmean"] = pred.abs().mean().item()
            fx[name + "_absmax"] = pred.abs().max().item()
            fx[name + "_rowsum"] = pred[0].sum(dim=2).numpy()  # [S, H, C]: every pixel enters a checksum
            print(name, tuple(pred.shape), "absmean", fx[name + "_absmean"], "absmax", fx[name + "_absmax"],
                  "conf mean", conf.mean().item())
    print("camera_params", fx["camera_params_steps"][-1])
    path = os.path.join(HERE, f"{case}_heads.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB")


if __name__ == "__main__":
    main()

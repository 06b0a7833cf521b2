"""Generate the Gaussian-path fixture from the REAL reference (run in the build container only).

    PYTHONPATH=/root/reference:/root/reference/submodules/gsplat python tests/golden/make_golden_gs.py

Config C1 (2 views 252x252, no priors).  The reference `VisualGeometryTransformer` (synthetic weights) produces
`token_list`; the reference `CameraHead`, `DPTHead(is_gsdpt=True)` and `GaussianSplatRenderer`, constructed exactly as
`WorldMirror._init_heads` does (worldmirror.py:66-118) and loaded strict=True with the synthetic head weights, produce
gs_depth / gs_depth_conf / the feature map, the raw Gaussian parameters, the splats before pruning (`prepare_splats`) and
after (`render`, whose gsplat rasterisation is stubbed: its images are dropped by the reference itself,
rasterization.py:239-246, and gsplat's CUDA extension cannot be built here).  Also stored: the reference `prune_gs` on the
seeded clustered splats of oracle/gs_oracle.py `synthetic_splats` at three voxel sizes, in full.
This fixture pins oracle/gs_oracle.py and is the target of the GPU parity tests of the Gaussian path.
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
from oracle import gs_oracle as G  # noqa: E402  (only for the seeded inputs)

STRIDE = 3
N_SYN = 12000


def main():
    from src.models.heads.camera_head import CameraHead  # reference
    from src.models.heads.dense_head import DPTHead  # reference
    from src.models.models.rasterization import GaussianSplatRenderer  # reference
    from src.models.models.visual_transformer import VisualGeometryTransformer  # reference

    n, h, w = 2, 252, 252
    agg = VisualGeometryTransformer(enable_cond=True).eval()
    agg.load_state_dict(W.synthetic_state_dict(seed=0), strict=True)
    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)

    def sub(p):
        return {k[len(p):]: v for k, v in hsd.items() if k.startswith(p)}

    cam = CameraHead(dim_in=2048).eval()
    cam.load_state_dict(sub("cam_head."), strict=True)
    gs_head = DPTHead(dim_in=2048, output_dim=2, patch_size=14, features=256, is_gsdpt=True, activation="exp+expp1").eval()
    gs_head.load_state_dict(sub("gs_head."), strict=True)
    rend = GaussianSplatRenderer(sh_degree=0, predict_offset=False, predict_residual_sh=True, enable_prune=True,
                                 voxel_size=0.002, using_gtcamera_splat=True, render_novel_views=True).eval()
    rend.load_state_dict(sub("gs_renderer."), strict=True)

    def no_raster(means, quats, scales, opacities, colors, viewmats, Ks, width, height, **kw):
        B, V = viewmats.shape[:2]
        z = torch.zeros(B, V, height, width, 1)
        return z.expand(-1, -1, -1, -1, 3), z, z

    rend.rasterizer.rasterize_batches = no_raster

    imgs = W.synthetic_images(n, h, w, seed=1)
    fx = {"n_views": n, "H": h, "W": w, "stride": STRIDE, "torch": torch.__version__, "n_syn": N_SYN}
    with torch.no_grad():
        token_list, psi = agg(imgs)
        cam_params = cam(token_list)[-1]  # [1, S, 9]
        gs_feat, gs_depth, gs_conf = gs_head([t.clone() for t in token_list], images=imgs, patch_start_idx=psi)
        fx["camera_params"] = cam_params[0].numpy()
        fx["gs_depth_sub"] = gs_depth[0, :, ::STRIDE, ::STRIDE, 0].numpy()
        fx["gs_conf_sub"] = gs_conf[0, :, ::STRIDE, ::STRIDE].numpy()
        fx["gs_depth_rowsum"] = gs_depth[0, ..., 0].sum(dim=2).numpy()
        fx["gs_feat_sub"] = gs_feat[0, :, :, ::12, ::12].numpy()
        fx["gs_feat_rowsum"] = gs_feat[0].sum(dim=3).numpy()  # [S, 128, H]
        fx["gs_feat_absmean"] = gs_feat.abs().mean().item()
        raw = rend.gs_head(gs_feat[0])  # [S, 12, H, W]
        fx["raw_sub"] = raw[:, :, ::STRIDE, ::STRIDE].permute(0, 2, 3, 1).numpy()
        fx["raw_absmean"] = raw.abs().mean(dim=(0, 2, 3)).numpy()
        preds = {"camera_params": cam_params, "gs_depth": gs_depth, "gs_depth_conf": gs_conf}
        views = {"img": imgs}
        splats = rend.prepare_splats(views, preds, imgs, raw, n, 0, position_from="gsdepth+predcamera")
        for k in ("means", "quats", "scales", "opacities", "sh", "weights"):
            a = splats[k][0].reshape(n * h * w, -1)
            fx["pre_" + k] = a[::7].numpy()
            fx["pre_" + k + "_sum"] = a.double().sum(dim=0).numpy()
        out = rend.render(gs_feats=gs_feat, images=imgs, predictions=dict(preds), views=views)["splats"]
        fx["pruned_K"] = out["means"][0].shape[0]
        for k in ("means", "quats", "scales", "opacities", "sh"):
            a = out[k][0].reshape(fx["pruned_K"], -1)
            fx["pruned_" + k] = a[::16].numpy()
            fx["pruned_" + k + "_sum"] = a.double().sum(dim=0).numpy()
        print("C1: gs_depth mean", gs_depth.mean().item(), "feat absmean", fx["gs_feat_absmean"], "raw absmean", fx["raw_absmean"],
              "N", n * h * w, "K", fx["pruned_K"])

        syn = G.synthetic_splats(N_SYN)
        tsyn = {k: torch.from_numpy(v)[None] for k, v in syn.items()}
        tsyn["sh"] = tsyn["sh"][:, :, None, :]
        for vs in G.PRUNE_VOXELS:
            o = rend.prune_gs(tsyn, voxel_size=vs)
            K = o["means"][0].shape[0]
            for k in ("means", "quats", "scales", "opacities", "sh"):
                fx[f"syn{vs}_{k}"] = o[k][0].reshape(K, -1).numpy()
            print("synthetic prune voxel", vs, "N", N_SYN, "K", K)
    path = os.path.join(HERE, "gs_c1.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB")


if __name__ == "__main__":
    main()

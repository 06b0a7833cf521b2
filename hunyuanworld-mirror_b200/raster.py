"""Forward Gaussian-splat rasterisation behind the reference's `Rasterizer` interface
(src/models/models/rasterization.py:17-95), SURVEY.md section 8f row 3.

The reference hands its splats to gsplat 1.5.3 (`rasterization(..., packed=True, rasterize_mode="classic",
render_mode="RGB+ED", sh_degree=0)`); inside `WorldMirror.forward` the images are then dropped
(rasterization.py:221-246), so `B200GaussianSplatRenderer.render` does not call this; it serves the novel-view / video
rendering callers (`infer.py:264`).  csrc/raster.cu restates gsplat's projection and tile binning; compositing follows gsplat's
arithmetic but streams packed per-intersection records through shared memory with bulk asynchronous copies.
There is no CPU path."""
from __future__ import annotations

import ctypes as C
from typing import Tuple

import torch

from . import capi

LAUNCHES = [0]


def _st():
    return torch.cuda.current_stream().cuda_stream


class B200Rasterizer:
    """`Rasterizer` (rasterization.py:17-95): rasterize_splats / rasterize_batches with the same arguments and return
    values (colors [V, H, W, 3], expected depths [V, H, W, 1], alphas [V, H, W, 1])."""

    def __init__(self, rasterization_mode="classic", packed=True, abs_grad=True, with_eval3d=False, camera_model="pinhole",
                 sparse_grad=False, distributed=False, grad_strategy=None):
        if rasterization_mode != "classic" or camera_model != "pinhole" or distributed or with_eval3d:
            raise NotImplementedError("B200Rasterizer covers the configuration WorldMirror uses: classic mode, pinhole cameras, one process")
        self.near_plane, self.far_plane, self.eps2d = 0.01, 1e10, 0.3
        self._ws = None

    @torch.no_grad()
    def rasterize_splats(self, means, quats, scales, opacities, colors, camtoworlds, Ks, width: int, height: int,
                         sh_degree=0, **kwargs) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        lib = capi.lib()
        if not means.is_cuda:
            raise RuntimeError("B200Rasterizer needs CUDA tensors (there is no CPU path)")
        if sh_degree not in (0, None) or colors.shape[-1] != 3 or colors.numel() != means.shape[0] * 3:
            raise NotImplementedError("B200Rasterizer renders degree-0 spherical harmonics [N, 1, 3]")
        dev = means.device
        f = lambda t: t.to(dev, torch.float32).contiguous()
        means, quats, scales, opacities, sh = f(means), f(quats), f(scales), f(opacities), f(colors).reshape(-1, 3)
        viewmats = torch.linalg.inv(f(camtoworlds)).contiguous()  # rasterization.py:46
        Ks = f(Ks)
        Cn, N = viewmats.shape[0], means.shape[0]
        total = Cn * N
        scan_b, rend_b = C.c_size_t(0), C.c_size_t(0)
        capi.check(lib.wm_raster_workspace_bytes(total, 0, C.byref(scan_b), C.byref(rend_b)))
        means2d = torch.empty(total, 2, dtype=torch.float32, device=dev)
        conics = torch.empty(total, 3, dtype=torch.float32, device=dev)
        depths = torch.empty(total, dtype=torch.float32, device=dev)
        radii = torch.empty(total, 2, dtype=torch.int32, device=dev)
        tiles = torch.empty(total, dtype=torch.int32, device=dev)
        cum = torch.empty(total, dtype=torch.int32, device=dev)
        scan_tmp = torch.empty(max(scan_b.value, 256), dtype=torch.uint8, device=dev)
        capi.check(lib.wm_raster_project(_st(), means.data_ptr(), quats.data_ptr(), scales.data_ptr(), opacities.data_ptr(),
                                         viewmats.data_ptr(), Ks.data_ptr(), Cn, N, width, height, self.near_plane, self.far_plane,
                                         self.eps2d, means2d.data_ptr(), conics.data_ptr(), depths.data_ptr(), radii.data_ptr(),
                                         tiles.data_ptr(), cum.data_ptr(), scan_tmp.data_ptr(), scan_tmp.numel()))
        n_isects = int(cum[-1].item()) & 0xFFFFFFFF  # one host read: sizes the sort buffers
        if n_isects >= 2 ** 31:
            raise RuntimeError(f"{n_isects} tile intersections exceed the 32-bit index range: render fewer views per call")
        capi.check(lib.wm_raster_workspace_bytes(total, n_isects, C.byref(scan_b), C.byref(rend_b)))
        if self._ws is None or self._ws.numel() < rend_b.value or self._ws.device != dev:
            self._ws = torch.empty(rend_b.value, dtype=torch.uint8, device=dev)
        tw, th = (width + 15) // 16, (height + 15) // 16
        offsets = torch.empty(Cn * tw * th + 1, dtype=torch.int32, device=dev)
        rgb = torch.empty(Cn, height, width, 3, dtype=torch.float32, device=dev)
        dep = torch.empty(Cn, height, width, 1, dtype=torch.float32, device=dev)
        alpha = torch.empty(Cn, height, width, 1, dtype=torch.float32, device=dev)
        capi.check(lib.wm_raster_render(_st(), means2d.data_ptr(), conics.data_ptr(), depths.data_ptr(), radii.data_ptr(),
                                        cum.data_ptr(), opacities.data_ptr(), sh.data_ptr(), Cn, N, width, height, n_isects,
                                        self._ws.data_ptr(), self._ws.numel(), offsets.data_ptr(), rgb.data_ptr(), dep.data_ptr(),
                                        alpha.data_ptr()))
        LAUNCHES[0] += 7
        self.last = {"means2d": means2d, "conics": conics, "depths": depths, "radii": radii, "tiles_per_gauss": tiles,
                     "n_isects": n_isects}
        return rgb, dep, alpha

    def rasterize_batches(self, means, quats, scales, opacities, colors, viewmats, Ks, width, height, **kwargs):
        """rasterization.py:68-94: lists of per-scene splats, viewmats [B, V, 4, 4] -> stacked [B, V, H, W, .]."""
        outs = [self.rasterize_splats(means[i], quats[i], scales[i], opacities[i], colors[i], viewmats[i], Ks[i], width, height, **kwargs)
                for i in range(len(means))]
        return tuple(torch.stack([o[k] for o in outs], dim=0) for k in range(3))

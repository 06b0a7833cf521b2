"""CPU oracle for the forward splat rasterisation.  TEST INFRASTRUCTURE ONLY (same import
rule as the other oracles).

numpy fp32 restatement of what the reference obtains from the vendored gsplat 1.5.3 through
`Rasterizer.rasterize_splats` (src/models/models/rasterization.py:29-66): pinhole cameras, classic mode, tile 16,
eps2d 0.3, near 0.01, far 1e10, SH degree 0, render_mode "RGB+ED".  The algorithm lives in a vendored dependency whose
CUDA extension cannot be built here (nested `glm` submodule absent), so it is restated from its sources:
  project    submodules/gsplat/gsplat/cuda/csrc/ProjectionEWA3DGSPacked.cu:58-190, include/Utils.cuh:142-222, 381-388, 498-537
  bin_tiles  csrc/IntersectTile.cu:24-110 (keys image | tile | depth bits, stable sort), :209 (offsets)
  composite  csrc/RasterizeToPixels3DGSFwd.cu:62-200; colours gsplat/rendering.py:509-525, expected depth :760-766

Parity status: PINNED against gsplat's own pure-torch statements, with one restated index list.  gsplat ships a
pure-torch statement of projection and tile binning (gsplat/cuda/_torch_impl.py: _quat_scale_to_covar_preci,
_fully_fused_projection :286, _isect_tiles :378, _isect_offset_encode :477) and of alpha compositing (`accumulate`
:506-598) that run on the CPU: tests/golden/make_golden_raster.py runs them on the seeded scene of `synthetic_scene`
below and tests/test_raster_oracle.py checks `project(opacity_aware=False)`, `bin_tiles` and `composite` against them.
Two things exist only in gsplat's CUDA kernels and are restated from their sources: the opacity-aware extent of the
projection, and the (gaussian, pixel) index list `accumulate` consumes (`rasterize_to_indices_in_range`,
csrc/RasterizeToIndices3DGS.cu:58-160: alpha-threshold skip and exclusive early stop), written out in the generating
script; `accumulate`'s two nerfacc helpers (not installed) are stubbed from their definitions there."""
from __future__ import annotations

from typing import Dict, Tuple

import numpy as np

f32 = np.float32
TILE = 16
ALPHA_MIN = f32(1.0 / 255.0)
C0 = f32(0.28209479177387814)


def quat_scale_to_covar(quats: np.ndarray, scales: np.ndarray) -> np.ndarray:
    """(w, x, y, z) quaternions, normalised, -> R S S R^T [N, 3, 3]."""
    q = quats.astype(f32)
    q = q / np.sqrt((q * q).sum(-1, keepdims=True, dtype=f32))
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                  2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                  2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], axis=-1).reshape(-1, 3, 3).astype(f32)
    M = R * scales.astype(f32)[:, None, :]
    return (M @ M.transpose(0, 2, 1)).astype(f32)


def project(means, quats, scales, opacities, viewmats, Ks, W: int, H: int, near=0.01, far=1e10, eps2d=0.3,
            opacity_aware: bool = True) -> Dict[str, np.ndarray]:
    """-> means2d [C,N,2], conics [C,N,3], depths [C,N], radii [C,N,2] int32 (0 = culled)."""
    means, viewmats, Ks = means.astype(f32), viewmats.astype(f32), Ks.astype(f32)
    cov = quat_scale_to_covar(quats, scales)
    R, t = viewmats[:, :3, :3], viewmats[:, :3, 3]
    mc = (np.einsum("cij,nj->cni", R, means) + t[:, None, :]).astype(f32)
    covc = np.einsum("cij,njk,clk->cnil", R, cov, R).astype(f32)
    fx, fy, cx, cy = (Ks[:, 0, 0, None], Ks[:, 1, 1, None], Ks[:, 0, 2, None], Ks[:, 1, 2, None])
    x, y, z = mc[..., 0], mc[..., 1], mc[..., 2]
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        tan_x, tan_y = f32(0.5) * W / fx, f32(0.5) * H / fy
        lim_xp, lim_xn = (W - cx) / fx + f32(0.3) * tan_x, cx / fx + f32(0.3) * tan_x
        lim_yp, lim_yn = (H - cy) / fy + f32(0.3) * tan_y, cy / fy + f32(0.3) * tan_y
        rz = f32(1.0) / z
        tx = z * np.minimum(lim_xp, np.maximum(-lim_xn, x * rz))
        ty = z * np.minimum(lim_yp, np.maximum(-lim_yn, y * rz))
        J = np.zeros(z.shape + (2, 3), dtype=f32)
        J[..., 0, 0], J[..., 0, 2] = fx * rz, -fx * tx * rz * rz
        J[..., 1, 1], J[..., 1, 2] = fy * rz, -fy * ty * rz * rz
        c2 = np.einsum("cnij,cnjk,cnlk->cnil", J, covc, J).astype(f32)
        m2 = np.stack([fx * x * rz + cx, fy * y * rz + cy], axis=-1).astype(f32)
        c00, c11, c01, c10 = c2[..., 0, 0] + f32(eps2d), c2[..., 1, 1] + f32(eps2d), c2[..., 0, 1], c2[..., 1, 0]
        det = c00 * c11 - c01 * c10
        valid = ~((z < f32(near)) | (z > f32(far))) & (det > 0)
        conics = np.stack([c11 / det, -c01 / det, c00 / det], axis=-1).astype(f32)
        o = opacities.astype(f32)[None, :]
        if opacity_aware:
            valid &= ~(o < ALPHA_MIN)
            extend = np.minimum(f32(3.33), np.sqrt(f32(2.0) * np.log(np.maximum(o / ALPHA_MIN, f32(1e-30)))))
        else:
            extend = f32(3.33)
        rx, ry = np.ceil(extend * np.sqrt(c00)), np.ceil(extend * np.sqrt(c11))
        valid &= ~((rx <= 0) & (ry <= 0))
        valid &= ~((m2[..., 0] + rx <= 0) | (m2[..., 0] - rx >= W) | (m2[..., 1] + ry <= 0) | (m2[..., 1] - ry >= H))
    radii = np.stack([np.where(valid, rx, 0), np.where(valid, ry, 0)], axis=-1)
    radii = np.nan_to_num(radii, nan=0.0, posinf=0.0, neginf=0.0).astype(np.int32)
    return {"means2d": m2, "conics": np.where(valid[..., None], conics, 0).astype(f32), "depths": z.astype(f32), "radii": radii,
            "valid": valid}


def tile_bounds(means2d, radii, tw: int, th: int):
    r = radii.astype(f32) / f32(TILE)
    c = means2d.astype(f32) / f32(TILE)
    lo = np.floor(c - r).astype(np.int64)
    hi = np.ceil(c + r).astype(np.int64)
    lim = np.array([tw, th])
    return np.clip(lo, 0, lim), np.clip(hi, 0, lim)


def bin_tiles(means2d, radii, depths, W: int, H: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """-> tiles_per_gauss [C,N], sorted isect_ids [M] int64, flatten_ids [M], offsets [C, th, tw]."""
    Cn, N = depths.shape
    tw, th = -(-W // TILE), -(-H // TILE)
    lo, hi = tile_bounds(means2d, radii, tw, th)
    ok = (radii > 0).all(-1)
    per = np.where(ok, (hi - lo).prod(-1), 0).astype(np.int64)
    tile_bits = int(tw * th).bit_length()
    ids, flat = [], []
    dbits = depths.astype(f32).view(np.uint32).astype(np.int64)
    for c in range(Cn):
        for g in np.nonzero(per[c])[0]:
            ys, xs = np.meshgrid(np.arange(lo[c, g, 1], hi[c, g, 1]), np.arange(lo[c, g, 0], hi[c, g, 0]), indexing="ij")
            tid = (ys * tw + xs).reshape(-1).astype(np.int64)
            ids.append((np.int64(c) << (32 + tile_bits)) | (tid << 32) | dbits[c, g])
            flat.append(np.full(tid.shape, c * N + g, dtype=np.int64))
    ids = np.concatenate(ids) if ids else np.zeros(0, np.int64)
    flat = np.concatenate(flat) if flat else np.zeros(0, np.int64)
    order = np.argsort(ids, kind="stable")
    ids, flat = ids[order], flat[order]
    tile_of = ((ids >> 32) >> tile_bits) * (tw * th) + ((ids >> 32) & ((1 << tile_bits) - 1))
    offsets = np.searchsorted(tile_of, np.arange(Cn * tw * th), side="left").reshape(Cn, th, tw)
    return per, ids, flat, offsets


def composite(proj: Dict[str, np.ndarray], opacities, sh, flat, offsets, W: int, H: int):
    """-> rgb [C,H,W,3], expected depth [C,H,W,1], alpha [C,H,W,1]."""
    Cn, th, tw = offsets.shape
    N = proj["depths"].shape[1]
    m2, con, dep = proj["means2d"].reshape(-1, 2), proj["conics"].reshape(-1, 3), proj["depths"].reshape(-1)
    col = np.maximum(C0 * sh.astype(f32).reshape(-1, 3) + f32(0.5), f32(0.0))
    ends = np.append(offsets.reshape(-1)[1:], len(flat))
    rgb = np.zeros((Cn, H, W, 3), f32)
    dout = np.zeros((Cn, H, W, 1), f32)
    aout = np.zeros((Cn, H, W, 1), f32)
    for c in range(Cn):
        for ty in range(th):
            for tx in range(tw):
                t = (c * th + ty) * tw + tx
                ys, xs = np.meshgrid(np.arange(ty * TILE, min((ty + 1) * TILE, H)), np.arange(tx * TILE, min((tx + 1) * TILE, W)), indexing="ij")
                px, py = xs.astype(f32) + f32(0.5), ys.astype(f32) + f32(0.5)
                T = np.ones(px.shape, f32)
                acc = np.zeros(px.shape + (4,), f32)
                done = np.zeros(px.shape, bool)
                for f in flat[offsets.reshape(-1)[t]:ends[t]]:
                    if done.all():
                        break
                    dx, dy = m2[f, 0] - px, m2[f, 1] - py
                    sigma = f32(0.5) * (con[f, 0] * dx * dx + con[f, 2] * dy * dy) + con[f, 1] * dx * dy
                    alpha = np.minimum(f32(0.999), opacities[f % N].astype(f32) * np.exp(-sigma).astype(f32))
                    use = ~done & ~(sigma < 0) & ~(alpha < ALPHA_MIN)
                    nT = T * (f32(1.0) - alpha)
                    stop = use & (nT <= f32(1e-4))
                    done |= stop
                    use &= ~stop
                    vis = np.where(use, alpha * T, f32(0.0)).astype(f32)
                    acc[..., :3] += col[f % N] * vis[..., None]
                    acc[..., 3] += dep[f] * vis
                    T = np.where(use, nT, T).astype(f32)
                a = f32(1.0) - T
                rgb[c, ys, xs] = acc[..., :3]
                aout[c, ys, xs, 0] = a
                dout[c, ys, xs, 0] = acc[..., 3] / np.maximum(a, f32(1e-10))
    return rgb, dout, aout


def rasterize(means, quats, scales, opacities, sh, camtoworlds, Ks, W: int, H: int):
    """Rasterizer.rasterize_splats (rasterization.py:29-66)."""
    viewmats = np.linalg.inv(camtoworlds.astype(np.float64)).astype(f32)
    p = project(means, quats, scales, opacities, viewmats, Ks, W, H)
    _, _, flat, offsets = bin_tiles(p["means2d"], p["radii"], p["depths"], W, H)
    return composite(p, opacities.astype(f32), sh, flat, offsets, W, H)


def synthetic_scene(n: int = 1500, cams: int = 2, W: int = 72, H: int = 52, seed: int = 11):
    """Seeded splats in front of `cams` pinhole cameras looking down +z from slightly different poses: a mix of tiny and
    screen-filling gaussians, some behind the camera, some off screen, some nearly transparent."""
    g = np.random.default_rng(seed)
    means = np.concatenate([g.uniform(-1.2, 1.2, size=(n, 2)), g.uniform(-0.5, 4.0, size=(n, 1))], axis=1)
    quats = g.normal(size=(n, 4))
    scales = np.exp(g.uniform(np.log(0.005), np.log(0.4), size=(n, 3)))
    opac = 1.0 / (1.0 + np.exp(-g.normal(0.5, 2.5, size=n)))
    sh = g.normal(0.0, 1.2, size=(n, 3))
    c2w = np.tile(np.eye(4), (cams, 1, 1))
    for c in range(cams):
        a = 0.15 * c
        c2w[c, :3, :3] = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])
        c2w[c, :3, 3] = [0.2 * c, -0.1 * c, -1.0]
    Ks = np.tile(np.array([[0.9 * W, 0, W / 2.0], [0, 0.9 * W, H / 2.0], [0, 0, 1.0]]), (cams, 1, 1))
    f = np.float32
    return {"means": means.astype(f), "quats": quats.astype(f), "scales": scales.astype(f), "opacities": opac.astype(f),
            "sh": sh.astype(f), "camtoworlds": c2w.astype(f), "Ks": Ks.astype(f), "W": W, "H": H}

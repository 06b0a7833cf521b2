"""CPU oracle for the depth-prior normalisation.  TEST INFRASTRUCTURE ONLY (same import
rule as the other oracles).

numpy restatement of `normalize_depth` (src/models/utils/priors.py:107-167) including torch.quantile's linear
interpolation (aten/src/ATen/native/Sorting.cpp quantile_compute: rank = q * (n - 1) in the input dtype, lerp between
the floor and ceil order statistics) in fp32.  Parity status: PINNED against the reference function itself:
tests/golden/make_golden_priors.py runs the real `normalize_depth` from /root/reference on the seeded inputs of
`synthetic_depth_cases` below and stores its outputs in tests/golden/priors_depth.npz."""
from __future__ import annotations

import numpy as np

f32 = np.float32


def _lerp(a: f32, b: f32, w: f32) -> f32:
    """aten/src/ATen/native/Lerp.h: weight < 0.5 ? a + w (b - a) : b - (b - a) (1 - w)."""
    d = f32(b - a)
    return f32(a + f32(w * d)) if abs(w) < 0.5 else f32(b - f32(d * f32(f32(1.0) - w)))


def _quantile(sorted_vals: np.ndarray, q: float) -> f32:
    n = sorted_vals.shape[0]
    rank = f32(f32(q) * f32(n - 1))
    lo = f32(np.floor(rank))
    w = f32(rank - lo)
    return _lerp(sorted_vals[int(lo)], sorted_vals[int(np.ceil(rank))], w)


def normalize_depth(depth: np.ndarray, eps: float = 1e-6, min_percentile: int = 1, max_percentile: int = 99) -> np.ndarray:
    """priors.py:107-167.  depth [B, S, H, W] fp32 -> same shape in [0, 1]."""
    B, S, H, W = depth.shape
    d = np.nan_to_num(depth.astype(f32).reshape(B * S, H * W), nan=0.0, posinf=1e6, neginf=0.0).astype(f32)  # :122
    out = np.empty_like(d)
    for i in range(d.shape[0]):
        flat = d[i]
        pos = flat[flat > 0]                                     # :130-134
        vals = np.sort(pos if pos.size > 0 else flat)
        if vals.size > 100:                                      # :137-144
            lo, hi = _quantile(vals, min_percentile / 100.0), _quantile(vals, max_percentile / 100.0)
        else:
            lo, hi = vals[0], vals[-1]
        if hi == lo:                                             # :147-148
            hi = f32(lo + f32(1.0))
        scale = abs(float(f32(hi - lo)))                         # :151-152: a Python double from .item()
        eps_val = max(eps, scale * eps)
        denom = f32(f32(hi - lo) + f32(eps_val))                 # the double is rounded to fp32 when added to the tensor
        out[i] = np.clip((flat - lo) / denom, 0.0, 1.0)          # :155-158
    return out.reshape(B, S, H, W)


def synthetic_depth_cases(seed: int = 0):
    """Seeded inputs covering the branches: ordinary maps, zeros mixed in, NaN / +-inf, fewer than 100 positive
    values (min/max branch), no positive value at all, a constant map (hi == lo)."""
    g = np.random.default_rng(seed)
    H, W = 70, 84
    a = g.uniform(0.5, 4.5, size=(H, W)).astype(f32)
    b = (g.gamma(2.0, 1.5, size=(H, W)) + 0.1).astype(f32)
    b[g.random((H, W)) < 0.3] = 0.0
    c = g.uniform(0.2, 9.0, size=(H, W)).astype(f32)
    c[0, :5] = np.nan
    c[1, :5] = np.inf
    c[2, :5] = -np.inf
    d = np.zeros((H, W), f32)
    d.reshape(-1)[g.choice(H * W, 60, replace=False)] = g.uniform(1.0, 2.0, 60).astype(f32)
    e = -g.uniform(0.0, 1.0, size=(H, W)).astype(f32)
    f = np.full((H, W), 2.5, f32)
    h = (g.uniform(0.0, 1.0, size=(H, W)) ** 8 * 100).astype(f32)  # heavy tail: percentiles far from min / max
    return np.stack([a, b, c, d, e, f, h])[None]  # [1, 7, H, W]


# --------------------------------------------------------------------------------------
# Pose / ray priors (worldmirror.py:228-249)
# --------------------------------------------------------------------------------------
def _rotmat_to_quat(m: np.ndarray) -> np.ndarray:
    """utils/rotation.py:41-97 for one 3x3 fp32 matrix -> quaternion xyzw with w >= 0."""
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = [f32(v) for v in m.reshape(9)]
    one = f32(1.0)
    t = np.array([one + m00 + m11 + m22, one + m00 - m11 - m22, one - m00 + m11 - m22, one - m00 - m11 + m22], dtype=f32)
    q_abs = np.where(t > 0, np.sqrt(np.maximum(t, f32(0))), f32(0)).astype(f32)
    cand = np.array([[q_abs[0] * q_abs[0], m21 - m12, m02 - m20, m10 - m01],
                     [m21 - m12, q_abs[1] * q_abs[1], m10 + m01, m02 + m20],
                     [m02 - m20, m10 + m01, q_abs[2] * q_abs[2], m12 + m21],
                     [m10 - m01, m20 + m02, m21 + m12, q_abs[3] * q_abs[3]]], dtype=f32)
    best = int(np.argmax(q_abs))
    q = (cand[best] / (f32(2.0) * np.maximum(q_abs[best], f32(0.1)))).astype(f32)  # r, i, j, k
    q = q[[1, 2, 3, 0]]
    return -q if q[3] < 0 else q


def pose_ray_priors(camera_pose: np.ndarray, intrinsics, H: int, W: int, padding: float = 0.1):
    """extract_priors' pose and ray branches (worldmirror.py:228-249) for one scene: camera_pose [S, 4, 4] (or [S, 3, 4]),
    intrinsics [S, 3, 3] | None -> (poses [S, 7], rays [S, 4] | None), fp32.
    normalize_poses: utils/priors.py:4-104; extrinsics_to_vector: utils/camera_utils.py:25-35."""
    ext = camera_pose[:, :3].astype(f32).copy()
    S = ext.shape[0]
    if not np.isfinite(ext).all():                                                     # priors.py:21-27
        ext = np.nan_to_num(ext, nan=0.0, posinf=1e6, neginf=-1e6).astype(f32)
    pos = ext[:, :3, 3]
    if S > 10:                                                                         # :56-63
        srt = np.sort(pos, axis=0)
        mn = np.array([_quantile(srt[:, a], 0.05) for a in range(3)], dtype=f32)
        mx = np.array([_quantile(srt[:, a], 0.95) for a in range(3)], dtype=f32)
    else:
        mn, mx = pos.min(axis=0), pos.max(axis=0)
    eps = np.maximum(f32(1e-6), np.abs(mx) * f32(1e-6)).astype(f32)                    # :69-73
    rng = np.maximum((mx - mn).astype(f32), eps)
    scale = f32(np.clip(rng.max(), f32(1e-6), f32(1e6)))                               # :77-78
    center = ((mn + mx) / f32(2.0)).astype(f32)                                        # :81
    actual = f32(scale / f32(1 - 2 * padding))                                         # :84
    t = np.clip(((pos - center) / actual + f32(0.5)).astype(f32), 0.0, 1.0)            # :85-88
    q = np.stack([_rotmat_to_quat(ext[s, :3, :3]) for s in range(S)])
    poses = np.concatenate([t, q], axis=1).astype(f32)
    rays = None
    if intrinsics is not None:                                                         # worldmirror.py:243-249
        k = intrinsics[:, :3, :3].astype(f32)
        rays = np.stack([k[:, 0, 0] / f32(W), k[:, 1, 1] / f32(H), k[:, 0, 2] / f32(W), k[:, 1, 2] / f32(H)], axis=1).astype(f32)
    return poses, rays


def synthetic_pose_cases(seed: int = 0):
    """[(camera_pose [S,4,4], intrinsics [S,3,3])]: 32 views (percentile branch), 5 views (min/max branch), 12 views with a
    NaN and an inf (nan_to_num branch), rotations with every quaternion component dominant once."""
    g = np.random.default_rng(seed)

    def rot(v):
        th = np.linalg.norm(v)
        k = v / th
        K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K

    cases = []
    for S, big in ((32, False), (5, True), (12, True)):
        pose = np.tile(np.eye(4, dtype=np.float64), (S, 1, 1))
        for s in range(S):
            ang = g.uniform(2.5, 3.1) if (big and s % 2) else g.uniform(0.01, 0.5)
            axis = g.normal(size=3)
            pose[s, :3, :3] = rot(axis / np.linalg.norm(axis) * ang)
            pose[s, :3, 3] = g.uniform(-1, 1, 3) * (5.0 if s == 3 else 1.0)
        intr = np.tile(np.array([[0.8 * 518, 0, 259.0], [0, 0.8 * 518, 259.0], [0, 0, 1.0]]), (S, 1, 1))
        intr[:, 0, 0] += g.uniform(-20, 20, S)
        intr[:, 1, 2] += g.uniform(-5, 5, S)
        cases.append((pose.astype(f32), intr.astype(f32)))
    cases[2][0][4, 1, 3] = np.nan
    cases[2][0][7, 0, 3] = np.inf
    return cases

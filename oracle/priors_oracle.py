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

"""CPU oracle for the image loader (second half).  TEST INFRASTRUCTURE ONLY (same import rule
as the other oracles).

numpy restatement of `prepare_images_to_tensor` (src/utils/inference_utils.py:187-248) from the decoded RGB array on:
`_calculate_resize_dims` (:159-171), `Image.resize(BICUBIC)`, `transforms.ToTensor()` (uint8 HWC -> fp32 CHW / 255), the
centre crop of the "crop" strategy (:222-225), `_apply_padding` with value 1.0 (:174-184) and the mixed-size padding
(:234-240).  The resize is Pillow's (a third-party dependency of the reference; Pillow 12.2.0 in this image), restated
from its published algorithm (src/libImaging/Resample.c: precompute_coeffs, normalize_coeffs_8bpc,
ImagingResampleHorizontal_8bpc / Vertical_8bpc): separable, horizontal pass first, bicubic a = -0.5 with the support
stretched by the down-scaling factor, coefficients normalised in double and rounded to 22-bit fixed point, uint8
intermediate image, result = clip8((2^21 + sum pixel * k) >> 22).  Integer arithmetic: the bar is bit-exact.

Parity status: PINNED against Pillow itself (tests/test_loader.py calls `PIL.Image.resize` on the seeded images below, and
the reference's own `prepare_images_to_tensor` generated tests/golden/loader_cases.npz through
tests/golden/make_golden_loader.py)."""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple

import numpy as np

PRECISION_BITS = 32 - 8 - 2


def bicubic(x: float) -> float:
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def coefficients(in_size: int, out_size: int) -> Tuple[int, np.ndarray, np.ndarray]:
    """precompute_coeffs + normalize_coeffs_8bpc for the whole-image box.  -> ksize, bounds [out, 2] (first, count),
    kk [out, ksize] int32."""
    scale = filterscale = in_size / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = 2.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        w = [bicubic((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        for x in range(xmax):
            v = w[x] / ww if ww != 0.0 else w[x]
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return ksize, bounds, kk


def _pass(img: np.ndarray, bounds: np.ndarray, kk: np.ndarray) -> np.ndarray:
    """One resampling pass along axis 0 of img [L, M, 3] uint8 -> [out, M, 3] uint8."""
    out = np.empty((bounds.shape[0],) + img.shape[1:], np.uint8)
    src = img.astype(np.int64)
    for i, (first, count) in enumerate(bounds):
        acc = (src[first:first + count] * kk[i, :count, None, None].astype(np.int64)).sum(axis=0) + (1 << (PRECISION_BITS - 1))
        out[i] = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return out


def resize_bicubic(img: np.ndarray, new_w: int, new_h: int) -> np.ndarray:
    """PIL.Image.resize((new_w, new_h), BICUBIC) on an RGB uint8 array [H, W, 3]."""
    H, W, _ = img.shape
    out = img
    if new_w != W:
        _, bh, kh = coefficients(W, new_w)
        need_rows = slice(0, H)
        if new_h != H:  # Pillow resamples only the rows the vertical pass reads; the result is the same
            pass
        out = _pass(out[need_rows].transpose(1, 0, 2), bh, kh).transpose(1, 0, 2)
    if new_h != H:
        _, bv, kv = coefficients(H, new_h)
        out = _pass(out, bv, kv)
    return np.ascontiguousarray(out)


def resize_dims(orig_w: int, orig_h: int, max_dim: int, strategy: str, patch: int = 14) -> Tuple[int, int]:
    """inference_utils.py:159-171 (Python round = banker's rounding)."""
    if strategy == "pad":
        if orig_w >= orig_h:
            new_w = max_dim
            new_h = round(orig_h * (new_w / orig_w) / patch) * patch
        else:
            new_h = max_dim
            new_w = round(orig_w * (new_h / orig_h) / patch) * patch
    else:
        new_w = max_dim
        new_h = round(orig_h * (new_w / orig_w) / patch) * patch
    return new_w, new_h


def _pad_square(t: np.ndarray, target: int) -> np.ndarray:
    """_apply_padding (:174-184): constant 1.0 up to target x target (never crops)."""
    h_pad, w_pad = target - t.shape[1], target - t.shape[2]
    if h_pad > 0 or w_pad > 0:
        top, left = max(h_pad, 0) // 2, max(w_pad, 0) // 2
        out = np.ones((3, t.shape[1] + max(h_pad, 0), t.shape[2] + max(w_pad, 0)), np.float32)
        out[:, top:top + t.shape[1], left:left + t.shape[2]] = t
        return out
    return t


def prepare_arrays(images: Sequence[np.ndarray], strategy: str = "crop", target: int = 518) -> np.ndarray:
    """prepare_images_to_tensor from decoded RGB uint8 arrays [H, W, 3] -> fp32 [1, N, 3, H', W']."""
    if not images:
        raise ValueError("At least 1 image is required")
    if strategy not in ("crop", "pad"):
        raise ValueError("Strategy must be 'crop' or 'pad'")
    tensors: List[np.ndarray] = []
    for img in images:
        h, w, _ = img.shape
        new_w, new_h = resize_dims(w, h, target, strategy)
        t = (resize_bicubic(img, new_w, new_h).astype(np.float32) / np.float32(255)).transpose(2, 0, 1)
        if strategy == "crop" and new_h > target:
            s = (new_h - target) // 2
            t = t[:, s:s + target, :]
        if strategy == "pad":
            t = _pad_square(t, target)
        tensors.append(t)
    dims = {(t.shape[1], t.shape[2]) for t in tensors}
    if len(dims) > 1:
        mh, mw = max(d[0] for d in dims), max(d[1] for d in dims)
        tensors = [_pad_square(t, max(mh, mw)) if (t.shape[1] != mh or t.shape[2] != mw) else t for t in tensors]
    return np.stack(tensors)[None]


def synthetic_photos(seed: int = 21):
    """Seeded uint8 'photos' of assorted sizes and aspect ratios (smooth gradients + texture + hard edges): down-scaling by
    large and small factors, up-scaling, portrait and landscape, sizes that trigger the crop and the pad branches."""
    g = np.random.default_rng(seed)
    out = []
    for (h, w) in ((480, 640), (1080, 1920), (700, 500), (300, 400), (518, 518), (1000, 333), (97, 131)):
        y, x = np.mgrid[0:h, 0:w]
        base = np.stack([127 + 120 * np.sin(x / 37.0 + y / 91.0), 127 + 120 * np.cos(x / 13.0), (x * 255.0 / w + y * 255.0 / h) / 2], axis=-1)
        img = base + g.normal(0, 25, size=(h, w, 3))
        img[h // 3:h // 3 + 7, :, :] = 255
        img[:, w // 2:w // 2 + 3, :] = 0
        out.append(np.clip(img, 0, 255).astype(np.uint8))
    return out

"""Pin oracle/raster_oracle.py (forward splat rasterisation) against gsplat's own pure-torch
statement of projection and tile binning (tests/golden/make_golden_raster.py -> raster_scene.npz), and check the
compositing restatement through properties (the reference has no CPU statement of it)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import raster_oracle as R


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(GOLDEN, "raster_scene.npz"))


def test_projection_and_binning_match_gsplat(fx):
    sc = R.synthetic_scene()
    W, H = sc["W"], sc["H"]
    np.testing.assert_allclose(R.quat_scale_to_covar(sc["quats"], sc["scales"]), fx["covars"], rtol=2e-5, atol=1e-8)
    p = R.project(sc["means"], sc["quats"], sc["scales"], sc["opacities"], fx["viewmats"], sc["Ks"], W, H, opacity_aware=False)
    vis = (fx["radii"] > 0).all(-1)
    assert np.array_equal((p["radii"] > 0).all(-1), vis)
    np.testing.assert_allclose(p["depths"], fx["depths"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(p["means2d"][vis], fx["means2d"][vis], rtol=1e-4, atol=2e-3)
    np.testing.assert_allclose(p["conics"][vis], fx["conics"][vis], rtol=2e-3, atol=1e-6)
    assert (np.abs(p["radii"][vis] - fx["radii"][vis]) <= 1).all() and (p["radii"][vis] != fx["radii"][vis]).mean() < 0.01
    # binning from the REFERENCE's projected values: identical keys, order and offsets
    per, ids, flat, offs = R.bin_tiles(fx["means2d"], fx["radii"], fx["depths"], W, H)
    assert np.array_equal(per, fx["tiles_per_gauss"])
    assert np.array_equal(ids, fx["isect_ids"]) and np.array_equal(flat, fx["flatten_ids"])
    assert np.array_equal(offs, fx["offsets"])


def test_compositing_properties():
    sc = R.synthetic_scene(n=400, cams=1, W=40, H=36, seed=3)
    args = (sc["means"], sc["quats"], sc["scales"], sc["opacities"], sc["sh"], sc["camtoworlds"], sc["Ks"], sc["W"], sc["H"])
    rgb, dep, alpha = R.rasterize(*args)
    assert rgb.shape == (1, 36, 40, 3) and dep.shape == (1, 36, 40, 1) and alpha.shape == (1, 36, 40, 1)
    assert (alpha >= 0).all() and (alpha <= 1 - 1e-4 + 1e-6).all() and alpha.max() > 0.5 and (rgb >= 0).all()
    covered = alpha[..., 0] > 1e-3
    assert (dep[..., 0][covered] > 0.01).all()
    # the result does not depend on the order the gaussians are listed in (depth sort decides), up to equal-depth ties
    perm = np.random.default_rng(0).permutation(400)
    rgb2, dep2, alpha2 = R.rasterize(*(a[perm] if i < 5 else a for i, a in enumerate(args)))
    np.testing.assert_allclose(rgb2, rgb, atol=1e-6)
    np.testing.assert_allclose(alpha2, alpha, atol=1e-6)
    # one opaque isotropic splat in front of the camera: alpha at its centre = min(0.999, o), colour = alpha * max(C0 sh + .5, 0)
    one = R.rasterize(np.array([[0, 0, 2.0]], np.float32), np.array([[1, 0, 0, 0]], np.float32), np.full((1, 3), 0.2, np.float32),
                      np.array([0.9], np.float32), np.array([[1.0, 0.0, -3.0]], np.float32), np.eye(4, dtype=np.float32)[None],
                      sc["Ks"], 40, 36)
    a = one[2][0, 18, 20, 0]
    assert 0.85 < a <= 0.9 + 1e-6
    np.testing.assert_allclose(one[0][0, 18, 20], a * np.array([R.C0 + 0.5, 0.5, 0.0]), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(one[1][0, 18, 20, 0], 2.0, rtol=1e-5)


def test_compositing_matches_gsplat_accumulate(fx):
    """`composite` against gsplat's own pure-torch alpha compositing (`accumulate`, _torch_impl.py:506-598) fed with the
    index list of `rasterize_to_indices_in_range` (tests/golden/make_golden_raster.py): colours, depth numerator and
    alpha of every pixel of the seeded scene, from the REFERENCE's projected values and tile lists.  fp32 vs fp32, same
    front-to-back order: agreement to rounding."""
    if "acc_renders" not in fx.files:
        pytest.skip("raster_scene.npz predates the compositing fixture")
    sc = R.synthetic_scene()
    W, H = sc["W"], sc["H"]
    proj = {"means2d": fx["means2d"], "conics": fx["conics"], "depths": fx["depths"]}
    rgb, dep, alpha = R.composite(proj, sc["opacities"], sc["sh"], fx["flatten_ids"], fx["offsets"], W, H)
    ref, ref_a = fx["acc_renders"], fx["acc_alphas"]
    assert ref.shape == (2, H, W, 4) and int(fx["acc_pairs"]) > 10000 and ref_a.max() > 0.9
    np.testing.assert_allclose(alpha, ref_a, atol=2e-6)
    np.testing.assert_allclose(rgb, ref[..., :3], atol=5e-6, rtol=1e-5)
    np.testing.assert_allclose(dep * np.maximum(alpha, 1e-10), ref[..., 3:], atol=2e-5, rtol=1e-5)  # "ED" = numerator / alpha

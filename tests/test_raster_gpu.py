"""GPU parity of the forward splat rasterisation (csrc/raster.cu, hunyuanworld-mirror_b200/raster.py) against the numpy
oracle (oracle/raster_oracle.py; its projection and binning are pinned against gsplat's torch statement), plus
size-independent properties on a scene of BASELINE size."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

pkg = importlib.import_module("hunyuanworld-mirror_b200")
raster = importlib.import_module("hunyuanworld-mirror_b200.raster")
from oracle import raster_oracle as R  # noqa: E402


def _render(sc, idx=None):
    keys = ("means", "quats", "scales", "opacities", "sh")
    t = {k: torch.from_numpy(sc[k] if idx is None or k not in keys else sc[k][idx]).cuda() for k in keys + ("camtoworlds", "Ks")}
    rz = raster.B200Rasterizer()
    out = rz.rasterize_splats(t["means"], t["quats"], t["scales"], t["opacities"], t["sh"][:, None, :], t["camtoworlds"], t["Ks"],
                              sc["W"], sc["H"], sh_degree=0)
    torch.cuda.synchronize()
    return rz, [o.cpu().numpy() for o in out]


@pytest.mark.parametrize("n,cams,W,H,seed", [(1500, 2, 72, 52, 11), (6000, 3, 150, 97, 5)])
def test_rasterizer_matches_oracle(n, cams, W, H, seed):
    sc = R.synthetic_scene(n, cams, W, H, seed)
    rz, (rgb, dep, alpha) = _render(sc)
    viewmats = np.linalg.inv(sc["camtoworlds"].astype(np.float64)).astype(np.float32)
    p = R.project(sc["means"], sc["quats"], sc["scales"], sc["opacities"], viewmats, sc["Ks"], W, H)
    last = {k: v.cpu().numpy() if torch.is_tensor(v) else v for k, v in rz.last.items()}
    radii = last["radii"].reshape(cams, n, 2)
    vis = (p["radii"] > 0).all(-1)
    gvis = (radii > 0).all(-1)
    assert (vis != gvis).mean() < 2e-3
    both = vis & gvis
    assert (np.abs(radii[both] - p["radii"][both]) <= 1).all() and (radii[both] != p["radii"][both]).mean() < 0.01
    np.testing.assert_allclose(last["means2d"].reshape(cams, n, 2)[both], p["means2d"][both], rtol=1e-4, atol=2e-3)
    np.testing.assert_allclose(last["conics"].reshape(cams, n, 3)[both], p["conics"][both], rtol=2e-3, atol=1e-6)
    np.testing.assert_allclose(last["depths"].reshape(cams, n), p["depths"], rtol=1e-5, atol=1e-6)
    ref_rgb, ref_dep, ref_alpha = R.rasterize(sc["means"], sc["quats"], sc["scales"], sc["opacities"], sc["sh"], sc["camtoworlds"], sc["Ks"], W, H)
    assert rgb.shape == ref_rgb.shape and dep.shape == ref_dep.shape and alpha.shape == ref_alpha.shape
    for name, got, ref in (("rgb", rgb, ref_rgb), ("alpha", alpha, ref_alpha)):
        d = np.abs(got - ref)
        bad = (d > 2e-3).mean()
        print(f"raster {n} splats {cams}x{H}x{W} {name}: mean abs diff {d.mean():.2e}, max {d.max():.2e}, pixels off by > 2e-3: {100 * bad:.3f} %")
        assert np.isfinite(got).all() and d.mean() < 2e-4 and bad < 5e-3
    cov = ref_alpha[..., 0] > 0.05
    dd = np.abs(dep - ref_dep)[..., 0][cov] / ref_dep[..., 0][cov]
    print(f"expected depth: mean rel diff {dd.mean():.2e}, pixels off by > 1e-3: {100 * (dd > 1e-3).mean():.3f} %")
    assert dd.mean() < 2e-4 and (dd > 1e-3).mean() < 5e-3


def test_rasterizer_properties_at_scale():
    """1 M splats, 4 views of 518 x 518: permutation invariance (the depth sort decides the order), alpha range, a blank
    view for a camera that looks away, and the batch interface."""
    g = np.random.default_rng(2)
    n, W, H = 1_000_000, 518, 518
    sc = R.synthetic_scene(16, 4, W, H, 1)
    sc.update({"means": np.concatenate([g.uniform(-1.5, 1.5, size=(n, 2)), g.uniform(0.5, 4.0, size=(n, 1))], axis=1).astype(np.float32),
               "quats": g.normal(size=(n, 4)).astype(np.float32), "scales": np.exp(g.uniform(np.log(2e-3), np.log(2e-2), size=(n, 3))).astype(np.float32),
               "opacities": g.uniform(0.0, 1.0, size=n).astype(np.float32), "sh": g.normal(size=(n, 3)).astype(np.float32)})
    sc["camtoworlds"][3, :3, :3] = np.diag([-1.0, 1.0, -1.0])  # looks down -z: nothing in front of it
    sc["camtoworlds"][3, :3, 3] = [0, 0, -1.0]
    rz, (rgb, dep, alpha) = _render(sc)
    print(f"1 M splats x 4 views: {rz.last['n_isects']} tile intersections")
    assert np.isfinite(rgb).all() and (alpha >= 0).all() and (alpha <= 1 - 1e-4 + 1e-6).all() and (rgb >= 0).all()
    assert alpha[:3].mean() > 0.5 and alpha[3].max() == 0 and rgb[3].max() == 0
    _, (rgb2, dep2, alpha2) = _render(sc, g.permutation(n))
    d = np.abs(rgb2 - rgb)
    assert d.mean() < 1e-5 and (d > 1e-3).mean() < 1e-3, (d.mean(), d.max())
    assert np.abs(alpha2 - alpha).mean() < 1e-5
    t = {k: torch.from_numpy(sc[k]).cuda() for k in ("means", "quats", "scales", "opacities", "sh", "camtoworlds", "Ks")}
    b = raster.B200Rasterizer().rasterize_batches([t["means"][:1000]] * 2, [t["quats"][:1000]] * 2, [t["scales"][:1000]] * 2,
                                                  [t["opacities"][:1000]] * 2, [t["sh"][:1000, None]] * 2, t["camtoworlds"][None].expand(2, -1, -1, -1),
                                                  t["Ks"][None].expand(2, -1, -1, -1), W, H, sh_degree=0)
    assert tuple(b[0].shape) == (2, 4, H, W, 3) and tuple(b[1].shape) == (2, 4, H, W, 1) and torch.equal(b[0][0], b[0][1])

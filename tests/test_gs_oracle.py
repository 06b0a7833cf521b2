"""Pin oracle/gs_oracle.py (Gaussian path) against the fixture the real reference modules
produced (tests/golden/make_golden_gs.py -> tests/golden/gs_c1.npz)."""
import importlib
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

W = importlib.import_module("hunyuanworld-mirror_b200.weights")
from oracle import gs_oracle as G  # noqa: E402

KEYS = ("means", "quats", "scales", "opacities", "sh")


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(GOLDEN, "gs_c1.npz"))


@pytest.fixture(scope="module")
def gs_sd():
    return W.synthetic_heads_state_dict(seed=0, gs=True)


@pytest.mark.parametrize("vs", G.PRUNE_VOXELS)
def test_prune_bit_equal_to_reference(fx, vs):
    """numpy restatement vs the reference's torch prune_gs on the same seeded splats: same voxels, same order, same bits
    (the quaternion norm is a 4-term reduction whose order torch does not promise: 2 ulp)."""
    out = G.prune_gs(G.synthetic_splats(int(fx["n_syn"])), voxel_size=vs)
    for k in KEYS:
        ref = fx[f"syn{vs}_{k}"]
        got = out[k].reshape(ref.shape)
        assert got.shape == ref.shape, (k, got.shape, ref.shape)
        if k == "quats":
            np.testing.assert_allclose(got, ref, rtol=3e-7, atol=1e-9)
        else:
            assert np.array_equal(got, ref), (k, np.abs(got - ref).max())


def test_gs_path_matches_reference(fx, gs_sd, c1_tokens):
    n, h, w, st = int(fx["n_views"]), int(fx["H"]), int(fx["W"]), int(fx["stride"])
    imgs = W.synthetic_images(n, h, w, seed=1)
    with torch.no_grad():
        feat, depth, conf = G.gs_dpt_head(c1_tokens, imgs, gs_sd)
        np.testing.assert_allclose(depth[0, :, ::st, ::st, 0].numpy(), fx["gs_depth_sub"], rtol=3e-3, atol=1e-4)
        np.testing.assert_allclose(conf[0, :, ::st, ::st].numpy(), fx["gs_conf_sub"], rtol=3e-3, atol=1e-4)
        scale = float(fx["gs_feat_absmean"])
        np.testing.assert_allclose(feat[:, :, ::12, ::12].numpy(), fx["gs_feat_sub"], rtol=3e-3, atol=3e-3 * scale)
        raw = G.gs_params(feat, gs_sd)
        np.testing.assert_allclose(raw[:, ::st, ::st].numpy(), fx["raw_sub"], rtol=3e-3, atol=3e-3)
        # prepare_splats from the oracle's own upstream values and the REFERENCE camera (fp32 vs fp32)
        sp = G.prepare_splats(raw, imgs[0], depth[0, ..., 0], torch.from_numpy(fx["camera_params"]))
    for k in KEYS + ("weights",):
        got = sp[k].reshape(n * h * w, -1)
        ref = fx["pre_" + k]
        tol = 5e-3 if k != "scales" else 2e-2  # scales = exp(raw ~ -7): a 3e-3 absolute error upstream is 3e-3 relative here
        np.testing.assert_allclose(got[::7].numpy(), ref, rtol=tol, atol=tol * float(np.abs(ref).mean()))
        np.testing.assert_allclose(got.double().sum(dim=0).numpy(), fx["pre_" + k + "_sum"], rtol=2e-3,
                                   atol=2e-3 * float(np.abs(ref).mean()) * got.shape[0])


def test_prune_of_reference_scale_scene(fx, gs_sd, c1_tokens):
    """prune at the model's voxel size on the oracle's C1 splats: the voxel count lands next to the reference's (the
    inputs differ in the last bits, so a few splats change voxel)."""
    n, h, w = int(fx["n_views"]), int(fx["H"]), int(fx["W"])
    imgs = W.synthetic_images(n, h, w, seed=1)
    with torch.no_grad():
        feat, depth, _ = G.gs_dpt_head(c1_tokens, imgs, gs_sd)
        sp = G.prepare_splats(G.gs_params(feat, gs_sd), imgs[0], depth[0, ..., 0], torch.from_numpy(fx["camera_params"]))
    out = G.prune_gs({k: v.numpy() for k, v in sp.items()}, voxel_size=0.002)
    K, Kref = out["means"].shape[0], int(fx["pruned_K"])
    assert abs(K - Kref) <= 0.002 * Kref, (K, Kref)
    for k in KEYS:
        ref = fx["pruned_" + k + "_sum"]
        got = out[k].reshape(K, -1).astype(np.float64).sum(axis=0)
        np.testing.assert_allclose(got, ref, rtol=5e-3, atol=5e-3 * float(np.abs(fx["pruned_" + k]).mean()) * K)


# ---------------------------------------------------------------------------------------------- confidence filter
@pytest.fixture(scope="module")
def cfx():
    return np.load(os.path.join(GOLDEN, "conf_filter.npz"))


def test_conf_filter_keeps_the_reference_set(cfx):
    """oracle `conf_filter_indices` vs the reference's apply_confidence_filter (rasterization.py:248-299) on seeded distinct
    confidences with masked (<= 1e-5) entries: the same kept set for every (percent, max_gaussians) case."""
    conf = cfx["conf"]
    for i, (p, mg) in enumerate(cfx["cases"]):
        kept = G.conf_filter_indices(conf, float(p), int(mg))
        assert kept.shape[0] == G.conf_filter_count(conf.shape[0], float(p), int(mg))
        assert np.array_equal(kept, cfx[f"kept{i}"]), (p, mg)


def test_conf_filter_then_prune_matches_reference(cfx):
    """filter -> prune_gs against the reference's result: same voxels; values to fp32 rounding (the reference accumulates a
    voxel in torch.topk's unspecified output order, the oracle in ascending index)."""
    syn = G.synthetic_splats(int(cfx["n_syn"]))
    out = G.prune_gs(G.apply_confidence_filter(syn, cfx["conf"], 30.0), voxel_size=0.01)
    for k in KEYS:
        ref = cfx["pruned_" + k]
        got = out[k].reshape(ref.shape[0], -1)
        assert got.shape == ref.shape, (k, got.shape, ref.shape)
        np.testing.assert_allclose(got, ref, rtol=2e-5, atol=1e-6)


def test_conf_filter_ties_and_masking():
    """Stated tie rule: equal confidences are taken in ascending index; everything <= 1e-5 ranks below every live value."""
    conf = np.array([2.0, 3.0, 2.0, 0.0, 2.0, 1e-6, 5.0, 2.0], dtype=np.float32)
    assert G.conf_filter_indices(conf, 50.0).tolist() == [0, 1, 2, 6]  # K = 4: 5, 3 and the first two 2.0
    assert G.conf_filter_indices(conf, 25.0).tolist() == [0, 1, 2, 4, 6, 7]  # K = 6: all live values
    assert G.conf_filter_indices(conf, 12.5).tolist() == [0, 1, 2, 3, 4, 6, 7]  # K = 7: the first masked one joins
    assert G.conf_filter_indices(conf, 0.0).tolist() == list(range(8))
    assert G.conf_filter_indices(conf, 30.0, max_gaussians=2).tolist() == [1, 6]
    # NaN ranks above everything, +inf included, as torch.topk has it
    c2 = np.array([1.5, np.nan, np.inf, 2.0, -np.nan, 0.0], dtype=np.float32)
    assert G.conf_filter_indices(c2, 50.0).tolist() == [1, 2, 4]
    t = torch.topk(torch.from_numpy(np.where(c2 <= 1e-5, -np.inf, c2)), 3, sorted=False).indices.sort().values.tolist()
    assert t == [1, 2, 4]


@pytest.mark.parametrize("seed", range(5))
def test_conf_filter_set_equals_torch_topk_on_distinct_values(seed):
    """The third-party call the reference makes (`torch.topk(conf, K, largest=True, sorted=False)`, rasterization.py:279) on
    random DISTINCT confidences with masked entries: the oracle keeps exactly that set, for random percentages."""
    g = np.random.default_rng(seed)
    n = int(g.integers(50, 3000))
    conf = g.permutation(np.linspace(0.5, 40.0, n)).astype(np.float32)
    conf[g.random(n) < 0.2] = 0.0
    p = float(g.uniform(0.0, 99.0))
    K = G.conf_filter_count(n, p)
    masked = torch.from_numpy(conf).masked_fill(torch.from_numpy(conf) <= 1e-5, float("-inf"))
    ref = torch.topk(masked, K, largest=True, sorted=False).indices.numpy()
    got = G.conf_filter_indices(conf, p)
    live = int((conf > 1e-5).sum())
    if K <= live:  # no choice among the equal -inf entries involved
        assert np.array_equal(np.sort(ref), got)
    else:  # all live entries kept; the rest are masked ones (which ones is unspecified in torch, lowest indices here)
        assert set(np.flatnonzero(conf > 1e-5)) <= set(got.tolist()) and got.shape[0] == K
        rest = np.setdiff1d(got, np.flatnonzero(conf > 1e-5))
        assert np.array_equal(rest, np.flatnonzero(conf <= 1e-5)[: K - live])

"""Depth-prior normalisation: the oracle against the fixture of the real reference
function (CPU), and the CUDA kernel against both (GPU)."""
import importlib
import os
import time

import numpy as np
import pytest
import torch

from conftest import GOLDEN

from oracle import priors_oracle as PO  # noqa: E402


def test_oracle_matches_reference_fixture():
    ref = np.load(os.path.join(GOLDEN, "priors_depth.npz"))["out"]
    got = PO.normalize_depth(PO.synthetic_depth_cases(seed=0))
    assert got.shape == ref.shape
    np.testing.assert_array_equal(got, ref)  # same fp32 operations in the same order: bit-exact


@pytest.mark.gpu
def test_cuda_normalize_depth_matches_reference_fixture():
    pkg = importlib.import_module("hunyuanworld-mirror_b200")
    ref = np.load(os.path.join(GOLDEN, "priors_depth.npz"))["out"]
    x = torch.from_numpy(PO.synthetic_depth_cases(seed=0)).cuda()
    got = pkg.priors.normalize_depth(x)
    torch.cuda.synchronize()
    assert tuple(got.shape) == ref.shape and got.dtype == torch.float32
    d = np.abs(got.cpu().numpy() - ref)
    print("normalize_depth vs reference fixture: max abs diff", d.max(), "bit-equal fraction", float((d == 0).mean()))
    assert d.max() <= 2.4e-7  # the percentiles are exact order statistics; the final division may differ by an ulp of 1.0


@pytest.mark.gpu
def test_cuda_normalize_depth_full_size_properties():
    """BASELINE config 3 size (32 views 518 x 518): against the oracle on two of the views, and size-independent
    properties on all: range [0, 1], monotone in the input, about 1 % of the positive pixels clamp at each end,
    invariance to a positive rescaling of a view."""
    pkg = importlib.import_module("hunyuanworld-mirror_b200")
    g = torch.Generator(device="cpu").manual_seed(3)
    x = torch.rand(1, 32, 518, 518, generator=g) * 4 + 0.5
    x[0, 5, :40] = 0.0
    xd = x.cuda()
    y = pkg.priors.normalize_depth(xd)
    t0 = time.perf_counter()
    y = pkg.priors.normalize_depth(xd)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"normalize_depth 32 x 518 x 518: {dt * 1e3:.3f} ms wall for one call")
    yc = y.cpu()
    ref = PO.normalize_depth(x[:, [0, 5]].numpy())
    assert np.abs(yc[:, [0, 5]].numpy() - ref).max() <= 2.4e-7
    assert float(yc.min()) >= 0.0 and float(yc.max()) <= 1.0
    v = x[0, 7].flatten()
    order = torch.argsort(v)
    assert bool((yc[0, 7].flatten()[order].diff() >= 0).all())
    frac0, frac1 = float((yc[0, 7] == 0).float().mean()), float((yc[0, 7] == 1).float().mean())
    assert 0.005 < frac0 < 0.015 and 0.005 < frac1 < 0.015
    y2 = pkg.priors.normalize_depth(xd * 8.0)  # power of two: exact rescaling
    assert float((y2 - y).abs().max()) <= 1e-6


def test_pose_oracle_matches_reference_fixture():
    fx = np.load(os.path.join(GOLDEN, "priors_pose.npz"))
    for i, (pose, intr) in enumerate(PO.synthetic_pose_cases(seed=0)):
        p, r = PO.pose_ray_priors(pose, intr, 518, 518)
        np.testing.assert_array_equal(p, fx[f"poses{i}"])
        np.testing.assert_array_equal(r, fx[f"rays{i}"])


@pytest.mark.gpu
def test_cuda_extract_priors_matches_reference_fixture():
    """`extract_priors` (worldmirror.py:218-251) on the GPU: poses, rays and depth of every fixture case."""
    pkg = importlib.import_module("hunyuanworld-mirror_b200")
    fx = np.load(os.path.join(GOLDEN, "priors_pose.npz"))
    dref = np.load(os.path.join(GOLDEN, "priors_depth.npz"))["out"]
    for i, (pose, intr) in enumerate(PO.synthetic_pose_cases(seed=0)):
        S = pose.shape[0]
        views = {"img": torch.zeros(1, S, 3, 518, 518, device="cuda"), "camera_pose": torch.from_numpy(pose)[None].cuda(),
                 "camera_intrinsics": torch.from_numpy(intr)[None].cuda()}
        depths, rays, poses = pkg.priors.extract_priors(views)
        torch.cuda.synchronize()
        assert depths is None and tuple(poses.shape) == (1, S, 7) and tuple(rays.shape) == (1, S, 4)
        dp = np.abs(poses[0].cpu().numpy() - fx[f"poses{i}"]).max()
        dr = np.abs(rays[0].cpu().numpy() - fx[f"rays{i}"]).max()
        print(f"pose case {i} ({S} views): poses max abs diff {dp:.2e}, rays {dr:.2e}")
        assert dp <= 1.2e-7 and dr == 0.0
    x = torch.from_numpy(PO.synthetic_depth_cases(seed=0)).cuda()
    d, r, p = pkg.priors.extract_priors({"img": torch.zeros(1, 7, 3, 70, 84, device="cuda"), "depthmap": x})
    assert r is None and p is None and np.abs(d.cpu().numpy() - dref).max() <= 2.4e-7

"""Pin oracle/heads_oracle.py (camera head + DPT heads) against the fixture the real
reference heads produced (tests/golden/make_golden_heads.py -> tests/golden/c1_heads.npz)."""
import importlib
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

W = importlib.import_module("hunyuanworld-mirror_b200.weights")
from oracle import aggregator_oracle as O  # noqa: E402
from oracle import heads_oracle as HO  # noqa: E402


@pytest.fixture(scope="module")
def heads_sd():
    return W.synthetic_heads_state_dict(seed=0)


def test_camera_head_matches_reference(heads_sd, c1_tokens):
    fx = np.load(os.path.join(GOLDEN, "c1_heads.npz"))
    # from the reference's own camera tokens: fp32 vs fp32
    with torch.no_grad():
        seq = HO.camera_head(torch.from_numpy(fx["cam_tokens"]), heads_sd)
    got = torch.stack(seq, 0).numpy()
    np.testing.assert_allclose(got, fx["camera_params_steps"], atol=2e-4, rtol=1e-4)
    # from the oracle aggregator's tokens (oracle end to end)
    with torch.no_grad():
        seq2 = HO.camera_head(c1_tokens[-1][:, :, 0], heads_sd)
    ang = HO.quat_angle(seq2[-1][..., 3:7], torch.from_numpy(fx["camera_params_steps"][-1][..., 3:7]))
    assert ang.max().item() < 1e-3, ang


@pytest.mark.parametrize("name,pre", [("depth", "depth_head."), ("pts3d", "pts_head."), ("normals", "norm_head.")])
def test_dpt_head_matches_reference(name, pre, heads_sd, c1_tokens):
    fx = np.load(os.path.join(GOLDEN, "c1_heads.npz"))
    st = int(fx["stride"])
    with torch.no_grad():
        pred, conf = HO.dpt_head(c1_tokens, 252, 252, heads_sd, pre, HO.HEAD_ACTIVATIONS[pre])
    assert tuple(pred.shape[:4]) == (1, 2, 252, 252) and tuple(conf.shape) == (1, 2, 252, 252)
    ref = fx[name + "_sub"]
    scale = float(fx[name + "_absmean"])
    np.testing.assert_allclose(pred[0, :, ::st, ::st].numpy(), ref, atol=2e-3 * scale, rtol=2e-3)
    np.testing.assert_allclose(conf[0, :, ::st, ::st].numpy(), fx[name + "_conf_sub"], rtol=2e-3, atol=1e-4)
    np.testing.assert_allclose(pred[0].sum(dim=2).numpy(), fx[name + "_rowsum"], atol=0.5 * scale, rtol=2e-3)

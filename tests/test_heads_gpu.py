"""GPU parity of the prediction-head kernels: each op against plain torch fp32 on the same
operands (convolutions run on the tensor cores at kind::f16 on fp16 operands, 10-bit mantissa like TF32, fp32 accumulation,
outputs rounded to fp16 = 2^-11 relative: tolerance 2^-10 of the output scale; same for the bandwidth kernels), and the whole
B200DPTHead against the fixture the REFERENCE heads produced (tests/golden/c1_heads.npz)."""
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
from oracle import heads_oracle as HO  # noqa: E402


def _rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return torch.randn(shape, generator=g) * scale


def _tf(x):
    """Operands exactly representable in the kernels' fp16 storage (10-bit mantissa)."""
    return heads.round_tf32(x)


def _st():
    return torch.cuda.current_stream().cuda_stream


def _close(got, ref, what, rel=2.0 ** -10):
    err = (got - ref).abs().max().item()
    tol = rel * ref.abs().max().item() + 1e-6
    print(f"{what}: max abs err {err:.3e} (ref max {ref.abs().max().item():.3f})")
    assert torch.isfinite(got).all() and err <= tol, (what, err, tol)


@pytest.mark.parametrize("n,H,Wd,cin,cout,taps,res,relu", [
    (2, 18, 18, 256, 256, 9, 2, 1),    # C1 grid, two residual inputs + ReLU (RCU1.conv2 of a fusion block)
    (1, 37, 37, 1024, 256, 9, 0, 1),   # layer3_rn at 518 px
    (2, 9, 11, 2048, 512, 1, 1, 0),    # projects[1] + positional table (broadcast residual), two Cout tiles
    (1, 40, 23, 256, 128, 9, 0, 0),    # output_conv1 shape class (Cout 128), ragged tiles in x and y
    (3, 5, 70, 128, 256, 1, 1, 1),     # image narrower than a tile row block
])
def test_conv2d_nhwc(n, H, Wd, cin, cout, taps, res, relu):
    k = 3 if taps == 9 else 1
    x = _tf(_rand((n, cin, H, Wd), 1))
    w = _tf(_rand((cout, cin, k, k), 2, 1.0 / (cin * taps) ** 0.5))
    b = _rand((cout,), 3, 0.1)
    bc = res == 1
    r1 = _rand((1 if bc else n, cout, H, Wd), 4) if res >= 1 else None
    r2 = _rand((n, cout, H, Wd), 5) if res >= 2 else None
    if r1 is not None and not bc:
        r1 = _tf(r1)
    if r2 is not None:
        r2 = _tf(r2)
    ref = F.conv2d(x, w, b, padding=k // 2)
    for r in (r1, r2):
        if r is not None:
            ref = ref + r
    if relu:
        ref = F.relu(ref)
    nhwc = lambda t, dt=torch.float16: None if t is None else t.permute(0, 2, 3, 1).contiguous().cuda().to(dt)
    xd, r1d, r2d = nhwc(x), nhwc(r1, torch.float32 if bc else torch.float16), nhwc(r2)
    wd = w.permute(0, 2, 3, 1).reshape(cout, taps * cin).contiguous().cuda().half()
    bd = b.cuda()
    out = torch.zeros(n, H, Wd, cout, dtype=torch.float16, device="cuda")
    capi.check(capi.lib().wm_conv2d_nhwc(_st(), xd.data_ptr(), n, H, Wd, cin, wd.data_ptr(), taps, cout, bd.data_ptr(),
                                         None if r1d is None else r1d.data_ptr(), int(bc), None if r2d is None else r2d.data_ptr(),
                                         relu, out.data_ptr()))
    torch.cuda.synchronize()
    _close(out.float().cpu().permute(0, 3, 1, 2), ref, f"conv {taps} taps {cin}->{cout} {H}x{Wd}")


@pytest.mark.parametrize("od,act", [(2, "exp"), (4, "inv_log"), (4, "norm")])
def test_conv2d_final(od, act):
    n, H, Wd, cin = 2, 21, 35, 128
    x = _tf(_rand((n, cin, H, Wd), 1))
    w1 = _tf(_rand((32, cin, 3, 3), 2, 1.0 / (cin * 9) ** 0.5))
    b1, w2, b2 = _rand((32,), 3, 0.1), _rand((od, 32, 1, 1), 4, 32 ** -0.5), _rand((od,), 5, 0.1)
    o = F.conv2d(F.relu(F.conv2d(x, w1, b1, padding=1)), w2, b2)
    pr, cr = HO.activate(o, act + "+expp1")
    xd = x.permute(0, 2, 3, 1).contiguous().cuda().half()
    w1d = w1.permute(0, 2, 3, 1).reshape(32, 9 * cin).contiguous().cuda().half()
    b1d, w2d, b2d = b1.cuda(), w2.reshape(od, 32).contiguous().cuda(), b2.cuda()
    pred = torch.zeros(n, H, Wd, od - 1, device="cuda")
    conf = torch.zeros(n, H, Wd, device="cuda")
    capi.check(capi.lib().wm_conv2d_final(_st(), xd.data_ptr(), n, H, Wd, cin, w1d.data_ptr(), b1d.data_ptr(), w2d.data_ptr(),
                                          b2d.data_ptr(), od, {"exp": 0, "inv_log": 1, "norm": 2}[act], pred.data_ptr(),
                                          conf.data_ptr()))
    torch.cuda.synchronize()
    _close(pred.cpu(), pr, f"final conv {act} pred", rel=1e-4)
    _close(conf.cpu(), cr, f"final conv {act} conf", rel=1e-4)


def test_bilinear_pixel_shuffle_im2col_ln():
    lib = capi.lib()
    n, Hi, Wi, c = 2, 19, 23, 256
    x = _tf(_rand((n, c, Hi, Wi), 1))
    xd = x.permute(0, 2, 3, 1).contiguous().cuda().half()
    for Ho, Wo, pos in ((37, 40, False), (38, 46, False), (60, 51, True)):
        ref = F.interpolate(x, size=(Ho, Wo), mode="bilinear", align_corners=True)
        pu = pv = None
        if pos:
            pu, pv = heads.uv_tables(Wo, Ho, c, Wo / Ho)
            ref = ref + HO.uv_pos_embed(Wo, Ho, c, Wo / Ho)[None]
            pu, pv = pu.cuda(), pv.cuda()
        out = torch.zeros(n, Ho, Wo, c, dtype=torch.float16, device="cuda")
        capi.check(lib.wm_bilinear_nhwc(_st(), xd.data_ptr(), n, Hi, Wi, c, Ho, Wo, None if pu is None else pu.data_ptr(),
                                        None if pv is None else pv.data_ptr(), out.data_ptr()))
        torch.cuda.synchronize()
        _close(out.float().cpu().permute(0, 3, 1, 2), ref, f"bilinear {Hi}x{Wi}->{Ho}x{Wo} pos={pos}")
    # ConvTranspose2d(k == stride) = 1x1 convolution to k*k*c channels + scatter
    for k in (4, 2):
        wt = _tf(_rand((c, c, k, k), 2, c ** -0.5))
        b = _rand((c,), 3, 0.1)
        ref = F.conv_transpose2d(x, wt, b, stride=k)
        wd = wt.permute(2, 3, 1, 0).reshape(k * k * c, c).contiguous().cuda().half()
        bd = b.repeat(k * k).contiguous().cuda()
        tmp = torch.zeros(n, Hi, Wi, k * k * c, dtype=torch.float16, device="cuda")
        capi.check(lib.wm_conv2d_nhwc(_st(), xd.data_ptr(), n, Hi, Wi, c, wd.data_ptr(), 1, k * k * c, bd.data_ptr(), None, 0, None, 0,
                                      tmp.data_ptr()))
        out = torch.zeros(n, Hi * k, Wi * k, c, dtype=torch.float16, device="cuda")
        capi.check(lib.wm_pixel_shuffle(_st(), tmp.data_ptr(), n, Hi, Wi, c, k, out.data_ptr()))
        torch.cuda.synchronize()
        _close(out.float().cpu().permute(0, 3, 1, 2), ref, f"conv_transpose k={k}")
    # stride-2 3x3 = im2col + 1x1 convolution
    w3 = _tf(_rand((c, c, 3, 3), 4, (9 * c) ** -0.5))
    b3 = _rand((c,), 5, 0.1)
    ref = F.conv2d(x, w3, b3, stride=2, padding=1)
    Ho, Wo = ref.shape[-2:]
    col = torch.zeros(n, Ho, Wo, 9 * c, dtype=torch.float16, device="cuda")
    capi.check(lib.wm_im2col_s2(_st(), xd.data_ptr(), n, Hi, Wi, c, col.data_ptr()))
    w3d = w3.permute(0, 2, 3, 1).reshape(c, 9 * c).contiguous().cuda().half()
    b3d = b3.cuda()
    out = torch.zeros(n, Ho, Wo, c, dtype=torch.float16, device="cuda")
    capi.check(lib.wm_conv2d_nhwc(_st(), col.data_ptr(), n, Ho, Wo, 9 * c, w3d.data_ptr(), 1, c, b3d.data_ptr(), None, 0, None, 0,
                                  out.data_ptr()))
    torch.cuda.synchronize()
    _close(out.float().cpu().permute(0, 3, 1, 2), ref, "conv3x3 stride 2")
    # LayerNorm(2048) of gathered rows
    S, P, psi, Cd = 3, 40, 7, 2048
    t = _rand((S, P, Cd), 6)
    lw, lb = 1 + _rand((Cd,), 7, 0.1), _rand((Cd,), 8, 0.1)
    ref = O.layer_norm(t[:, psi:], lw, lb, 1e-5)
    td, lwd, lbd = t.cuda(), lw.cuda(), lb.cuda()
    y = torch.zeros(S, P - psi, Cd, dtype=torch.float16, device="cuda")
    capi.check(lib.wm_layernorm_rows(_st(), td.data_ptr(), lwd.data_ptr(), lbd.data_ptr(), 1e-5, S * (P - psi), P - psi, P, psi,
                                     Cd, y.data_ptr()))
    torch.cuda.synchronize()
    _close(y.float().cpu(), ref, "layernorm rows 2048")


@pytest.mark.parametrize("name,pre,od", [("depth", "depth_head.", 2), ("pts3d", "pts_head.", 4), ("normals", "norm_head.", 4)])
def test_dpt_head_matches_reference_golden(name, pre, od, state_dict_full):
    """B200DPTHead on the fp32 oracle aggregator's tokens (pinned to the reference) against what the REFERENCE DPTHead
    produced from the reference tokens.  Gate: rel. Frobenius <= 2e-3 (tf32 operands, fp32 accumulation and storage)."""
    fx = np.load(os.path.join(GOLDEN, "c1_heads.npz"))
    st = int(fx["stride"])
    hsd = W.synthetic_heads_state_dict(seed=0)
    imgs = W.synthetic_images(2, 252, 252, seed=1)
    with torch.no_grad():
        toks, psi = O.aggregator_forward(imgs, state_dict_full)
    head = heads.B200DPTHead(dim_in=2048, output_dim=od, activation=HO.HEAD_ACTIVATIONS[pre])
    head.load_state_dict({k[len(pre):]: v for k, v in hsd.items() if k.startswith(pre)})
    pred, conf = head([t.cuda() for t in toks], imgs.cuda(), psi)
    torch.cuda.synchronize()
    assert tuple(pred.shape) == (1, 2, 252, 252, od - 1) and tuple(conf.shape) == (1, 2, 252, 252)
    got, ref = pred[0, :, ::st, ::st].cpu().numpy(), fx[name + "_sub"]
    gc, rc = conf[0, :, ::st, ::st].cpu().numpy(), fx[name + "_conf_sub"]
    e = np.linalg.norm(got - ref) / np.linalg.norm(ref)
    ec = np.linalg.norm(gc - rc) / np.linalg.norm(rc)
    rs = np.abs(pred[0].sum(dim=2).cpu().numpy() - fx[name + "_rowsum"]).max() / (252 * float(fx[name + "_absmean"]))
    print(f"{name}: rel fro {e:.3e}, conf rel fro {ec:.3e}, row-sum dev {rs:.3e}")
    assert np.isfinite(got).all() and e <= 2e-3 and ec <= 2e-3 and rs <= 5e-3


def test_camera_head_matches_reference_golden():
    """B200CameraHead (fp32 SIMT) on the reference's own camera tokens against the reference CameraHead's four
    refinement steps: fp32 vs fp32, different summation order only."""
    fx = np.load(os.path.join(GOLDEN, "c1_heads.npz"))
    hsd = W.synthetic_heads_state_dict(seed=0)
    cam = heads.B200CameraHead(dim_in=2048)
    cam.load_state_dict({k[len("cam_head."):]: v for k, v in hsd.items() if k.startswith("cam_head.")})
    S = fx["cam_tokens"].shape[1]
    feat = torch.zeros(1, S, 5, 2048)  # the camera token is row 0 of every view; other rows must not matter
    feat[:, :, 0] = torch.from_numpy(fx["cam_tokens"])
    feat[:, :, 1:] = 7.0
    seq = cam([feat.cuda()])
    torch.cuda.synchronize()
    got = torch.stack(seq, 0).cpu().numpy()
    np.testing.assert_allclose(got, fx["camera_params_steps"], atol=3e-4, rtol=2e-4)
    ang = HO.quat_angle(torch.from_numpy(got[-1][..., 3:7]), torch.from_numpy(fx["camera_params_steps"][-1][..., 3:7])).max().item()
    print(f"camera head vs reference: max abs {np.abs(got - fx['camera_params_steps']).max():.3e}, rotation {ang:.3e} rad")
    assert ang < 1e-4


def test_camera_head_many_views_matches_oracle():
    """S = 21 views (three row tiles of the skinny GEMM, a ragged last one) against the fp32 oracle."""
    hsd = W.synthetic_heads_state_dict(seed=0)
    cam = heads.B200CameraHead(dim_in=2048)
    cam.load_state_dict({k[len("cam_head."):]: v for k, v in hsd.items() if k.startswith("cam_head.")})
    toks = _rand((1, 21, 2048), 11)
    with torch.no_grad():
        ref = HO.camera_head(toks, hsd)
    seq = cam([toks[:, :, None].contiguous().cuda()])
    torch.cuda.synchronize()
    for a, b in zip(seq, ref):
        np.testing.assert_allclose(a.cpu().numpy(), b.numpy(), atol=3e-4, rtol=2e-4)


def test_worldmirror_forward_end_to_end(state_dict_full):
    """The whole B200 path (bf16 aggregator, fp32 camera head, tf32 DPT heads) against the REFERENCE aggregator +
    REFERENCE heads (tests/golden/c1_heads.npz), BASELINE.json config 1 shapes.  Gates as in
    tests/test_parity_gpu.py::test_north_star_tolerances_end_to_end."""
    fx = np.load(os.path.join(GOLDEN, "c1_heads.npz"))
    st = int(fx["stride"])
    model = heads.B200WorldMirror.from_state_dicts(state_dict_full, W.synthetic_heads_state_dict(seed=0))
    preds = model({"img": W.synthetic_images(2, 252, 252, seed=1).cuda()})
    torch.cuda.synchronize()
    for name in ("depth", "pts3d", "normals"):
        got, ref = preds[name][0, :, ::st, ::st].cpu().numpy(), fx[name + "_sub"]
        e = np.linalg.norm(got - ref) / np.linalg.norm(ref)
        eps = 1e-2 * np.median(np.abs(ref))
        mr = (np.abs(got - ref) / np.maximum(np.abs(ref), eps)).max()
        print(f"e2e {name}: rel fro {e:.3e} max-rel {mr:.3e}")
        assert np.isfinite(got).all() and e <= 1e-2
    cam_ref = torch.from_numpy(fx["camera_params_steps"][-1])
    ang = HO.quat_angle(preds["camera_params"].cpu()[..., 3:7], cam_ref[..., 3:7]).max().item()
    print(f"e2e camera rotation {ang:.3e} rad")
    assert ang <= 1e-3


def test_dpt_head_518_matches_oracle():
    """BASELINE.json's image size (518 x 518: 37 x 37 patch grid, maps of 19 / 37 / 74 / 148 / 296 / 518 pixels, every
    border-tile case of the convolution kernel) on one view of synthetic tokens against the fp32 oracle head."""
    hsd = W.synthetic_heads_state_dict(seed=0)
    pre = "pts_head."
    P = 7 + 37 * 37
    toks = [_rand((1, 1, P, 2048), 20 + i, 0.8) for i in range(4)]
    with torch.no_grad():
        ref_p, ref_c = HO.dpt_head(toks, 518, 518, hsd, pre, "inv_log+expp1")
    head = heads.B200DPTHead(dim_in=2048, output_dim=4, activation="inv_log+expp1")
    head.load_state_dict({k[len(pre):]: v for k, v in hsd.items() if k.startswith(pre)})
    pred, conf = head([t.cuda() for t in toks], torch.zeros(1, 1, 3, 518, 518, device="cuda"), 7)
    torch.cuda.synchronize()
    e = ((pred.cpu() - ref_p).norm() / ref_p.norm()).item()
    ec = ((conf.cpu() - ref_c).norm() / ref_c.norm()).item()
    border = (pred.cpu() - ref_p)[0, 0, [0, -1]].abs().max().item() / ref_p.abs().max().item()
    print(f"DPT 518: rel fro {e:.3e}, conf {ec:.3e}, border rows max err / max {border:.3e}")
    assert torch.isfinite(pred).all() and e <= 2e-3 and ec <= 2e-3 and border <= 5e-3


def test_single_view_forward_matches_oracle_heads(state_dict_full):
    """Edge case S = 1 (one view, rectangular 42 x 70): the whole `B200WorldMirror.forward` including the Gaussian branch.
    The heads are checked against the fp32 oracle heads run on the B200 aggregator's OWN tokens (the forward is
    deterministic), so the gates are the heads' own: dense outputs 2e-3 rel. Frobenius, camera 3e-4; the splats are the
    oracle's voxel merge of this forward's un-pruned splats, bit for bit (quaternions 2 ulp)."""
    from oracle import gs_oracle as G

    hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    model = heads.B200WorldMirror.from_state_dicts(state_dict_full, hsd)
    imgs = W.synthetic_images(1, 42, 70, seed=3).cuda()
    preds = model({"img": imgs})
    token_list, psi = model.visual_geometry_transformer(imgs)
    torch.cuda.synchronize()
    assert psi == 7 and token_list[0].shape == (1, 1, 7 + 3 * 5, 2048)
    with torch.no_grad():
        ref = HO.all_heads([t.cpu() for t in token_list], 42, 70, hsd)
    for name in ("depth", "pts3d", "normals", "depth_conf", "pts3d_conf", "normals_conf"):
        got, r = preds[name].cpu().numpy(), ref[name].numpy()
        assert got.shape == r.shape, (name, got.shape, r.shape)
        e = np.linalg.norm(got - r) / np.linalg.norm(r)
        print(f"single view {name}: rel fro {e:.3e}")
        assert np.isfinite(got).all() and e <= 2e-3
    np.testing.assert_allclose(preds["camera_params"].cpu().numpy(), ref["camera_params"].numpy(), atol=3e-4, rtol=2e-4)
    K = preds["splats"]["means"][0].shape[0]
    model.gs_renderer.enable_prune = False
    raw = model({"img": imgs})["splats"]
    torch.cuda.synchronize()
    keys = ("means", "quats", "scales", "opacities", "sh")
    mref = G.prune_gs({k: raw[k][0].reshape(raw[k].shape[1], -1).squeeze(-1).cpu().numpy() for k in keys + ("weights",)}, 0.002)
    assert mref["means"].shape[0] == K and 0 < K <= 42 * 70
    for k in keys:
        g = preds["splats"][k][0].reshape(K, -1).cpu().numpy()
        if k == "quats":
            np.testing.assert_allclose(g, mref[k], rtol=3e-7, atol=1e-9)
        else:
            assert np.array_equal(g, mref[k].reshape(K, -1)), k


@pytest.mark.parametrize("tag,flags", [("nopriors", (0, 0, 0)), ("priors", (1, 1, 1))])
def test_worldmirror_predictions_match_reference_forward(tag, flags, state_dict_full):
    """`B200WorldMirror.forward` against the dictionary the REFERENCE `WorldMirror.forward` (worldmirror.py:120-220,
    enable_gs=False) returned on config C1 (tests/golden/make_golden_preds.py): the same key set, the same shapes, the
    camera-to-world poses and intrinsics (worldmirror.py:161-175) and the dense predictions on a stride-6 lattice.
    Gates: dense rel. Frobenius <= 1e-2 (north_star), camera parameters 2e-3 absolute, rotation <= 1e-3 rad,
    intrinsics 0.5 % of the focal length."""
    fx = np.load(os.path.join(GOLDEN, "c1_preds.npz"))
    n, h, w, st = int(fx["n_views"]), int(fx["H"]), int(fx["W"]), int(fx["stride"])
    model = heads.B200WorldMirror.from_state_dicts(state_dict_full, W.synthetic_heads_state_dict(seed=0))
    views = {"img": W.synthetic_images(n, h, w, seed=1).cuda()}
    if sum(flags):
        pose, depth, intr = W.synthetic_scene(n, h, w, seed=2)
        views.update(camera_pose=pose.cuda(), depthmap=depth.cuda(), camera_intrinsics=intr.cuda())
    preds = model(views, cond_flags=flags)
    torch.cuda.synchronize()
    ref_keys = [str(k) for k in fx[tag + "_keys"]]
    assert sorted(preds.keys()) == ref_keys, (sorted(preds.keys()), ref_keys)
    for k, shp in zip(ref_keys, fx[tag + "_shapes"]):
        assert str(tuple(preds[k].shape)) == str(shp), (k, tuple(preds[k].shape), shp)
    for name in ("depth", "depth_conf", "pts3d", "pts3d_conf", "normals", "normals_conf"):
        got, ref = preds[name][0, :, ::st, ::st].cpu().numpy(), fx[f"{tag}_{name}_sub"]
        e = np.linalg.norm(got - ref) / np.linalg.norm(ref)
        print(f"{tag} {name}: rel fro {e:.3e}")
        assert np.isfinite(got).all() and e <= 1e-2, (name, e)
    cp, cp_ref = preds["camera_params"].cpu(), torch.from_numpy(fx[tag + "_camera_params"])
    assert (cp - cp_ref).abs().max().item() <= 2e-3
    poses, poses_ref = preds["camera_poses"].cpu().numpy(), fx[tag + "_camera_poses"]
    intrs, intrs_ref = preds["camera_intrs"].cpu().numpy(), fx[tag + "_camera_intrs"]
    ang = HO.quat_angle(cp[..., 3:7], cp_ref[..., 3:7]).max().item()
    print(f"{tag} cameras: |params| {(cp - cp_ref).abs().max().item():.3e}, rotation {ang:.3e} rad, "
          f"|poses| {np.abs(poses - poses_ref).max():.3e}, |K| {np.abs(intrs - intrs_ref).max():.3e}")
    assert ang <= 1e-3
    np.testing.assert_allclose(poses, poses_ref, atol=3e-3)
    np.testing.assert_allclose(intrs, intrs_ref, rtol=5e-3, atol=5e-3 * float(intrs_ref[..., 0, 0].max()))
    np.testing.assert_array_equal(poses[..., 3, :], poses_ref[..., 3, :])


def test_worldmirror_forward_c2_matches_reference_golden(state_dict_full):
    """The configuration bench.py TIMES (BASELINE.json configs[1]: 8 views 518 x 518, no priors): the whole B200 forward
    (bf16 aggregator with the default attention variants, fp32 camera head, fp16-operand DPT heads) against the REFERENCE
    aggregator + REFERENCE heads run in fp32 on the same inputs (tests/golden/c2_heads.npz, stride-7 lattice + row sums
    of every pixel).  Gates: north_star's 1e-2 rel. Frobenius on every dense map, 1e-3 rad on the camera rotation."""
    fx = np.load(os.path.join(GOLDEN, "c2_heads.npz"))
    n, h, w, st = int(fx["n_views"]), int(fx["H"]), int(fx["W"]), int(fx["stride"])
    model = heads.B200WorldMirror.from_state_dicts(state_dict_full, W.synthetic_heads_state_dict(seed=0))
    preds = model({"img": W.synthetic_images(n, h, w, seed=1).cuda()})
    torch.cuda.synchronize()
    for name in ("depth", "pts3d", "normals"):
        got, ref = preds[name][0, :, ::st, ::st].cpu().numpy(), fx[name + "_sub"]
        gc, rc = preds[name + "_conf"][0, :, ::st, ::st].cpu().numpy(), fx[name + "_conf_sub"]
        e = np.linalg.norm(got - ref) / np.linalg.norm(ref)
        ec = np.linalg.norm(gc - rc) / np.linalg.norm(rc)
        rs = np.abs(preds[name][0].sum(dim=2).cpu().numpy() - fx[name + "_rowsum"]).max() / (w * float(fx[name + "_absmean"]))
        print(f"C2 e2e {name}: rel fro {e:.3e}, conf {ec:.3e}, row-sum dev {rs:.3e}")
        assert np.isfinite(got).all() and e <= 1e-2 and ec <= 1e-2 and rs <= 2e-2, (name, e, ec, rs)
    cam_ref = torch.from_numpy(fx["camera_params_steps"][-1])
    cp = preds["camera_params"].cpu()
    ang = HO.quat_angle(cp[..., 3:7], cam_ref[..., 3:7]).max().item()
    print(f"C2 e2e camera: |params| {(cp - cam_ref).abs().max().item():.3e}, rotation {ang:.3e} rad")
    assert ang <= 1e-3 and (cp - cam_ref).abs().max().item() <= 5e-3


@pytest.mark.parametrize("scale", [1e3, 1e6])
def test_fp16_range_guard(scale, c1_tokens):
    """The DPT heads keep their maps in fp16 (the reference: fp32 / bf16-autocast, dense_head.py:164-251).  With the
    normalised tokens scaled by `scale` (norm.weight and norm.bias of the head) the head must EITHER stay in parity with
    the fp32 oracle on the same weights OR report the saturation (`fp16_saturation()` > 0, which `B200WorldMirror.forward`
    turns into Fp16RangeError); 1e6 is far outside the range and must be reported.  The normals head is used: its
    output x / |x| stays finite at any scale (the confidence 1 + exp(.) is compared where the oracle's is finite)."""
    toks, psi = c1_tokens, 7
    hsd = dict(W.synthetic_heads_state_dict(seed=0))
    pre = "norm_head."
    hsd[pre + "norm.weight"] = hsd[pre + "norm.weight"] * scale
    hsd[pre + "norm.bias"] = hsd[pre + "norm.bias"] * scale
    head = heads.B200DPTHead(dim_in=2048, output_dim=4, activation="norm+expp1")
    try:
        head.load_state_dict({k[len(pre):]: v for k, v in hsd.items() if k.startswith(pre)})
    except heads.Fp16RangeError as e:  # the folded projection weights themselves leave the fp16 range: reported at load time
        print(f"scale {scale:g}: {e}")
        assert scale >= 1e6
        return
    imgs = W.synthetic_images(2, 252, 252, seed=1)
    heads.fp16_saturation(reset=True)
    pred, conf = head([t.cuda() for t in toks], imgs.cuda(), psi)
    sat = heads.fp16_saturation(reset=True)
    print(f"scale {scale:g}: {sat} saturated (pixel, 32-channel) entries")
    if scale >= 1e6:
        assert sat > 0
    if sat == 0:
        with torch.no_grad():
            ref_p, ref_c = HO.dpt_head(toks, 252, 252, hsd, pre, "norm+expp1")
        assert torch.isfinite(ref_p).all() and torch.isfinite(pred).all()
        e = ((pred.cpu() - ref_p).norm() / ref_p.norm()).item()
        print(f"  in range: normals rel fro {e:.3e}")
        assert e <= 5e-3
    else:
        class _Agg:  # a stand-in aggregator that returns the fixture's tokens: only the guard of the forward is under test
            world = 1

            def __call__(self, imgs, *a, **k):
                return [t.cuda() for t in toks], psi
        model = heads.B200WorldMirror(_Agg(), norm_head=head)
        with pytest.raises(heads.Fp16RangeError):
            model({"img": imgs.cuda()})  # no Gaussian branch: the count is checked by the NEXT forward
            torch.cuda.synchronize()
            model({"img": imgs.cuda()})
        heads.fp16_saturation(reset=True)

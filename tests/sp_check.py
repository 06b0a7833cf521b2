"""Run under torchrun on >= 2 B200s (tests/test_sp_gpu.py launches it): sequence-parallel forward with the
peer-memory K|V exchange against (a) the ncclAllGather path and (b) a single-GPU forward of all views."""
import importlib
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    pkg = importlib.import_module("hunyuanworld-mirror_b200")
    W, capi = pkg.weights, pkg.capi
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    depth, dino_depth, per, H, Wd = 3, 1, 2, 126, 154  # P = 7 + 9*11 = 106 rows/view: shards of 212 rows, not tile-aligned
    sd = W.synthetic_state_dict(seed=0, depth=depth, dino_depth=dino_depth)
    m = pkg.B200VisualGeometryTransformer(depth=depth, dino_depth=dino_depth, intermediate_idxs=(0, 2))
    m.load_state_dict(sd)
    m.init_sequence_parallel()
    imgs = W.synthetic_images(per * world, H, Wd, seed=1)
    pri = W.synthetic_priors(per * world, H, Wd, seed=2)
    sl = slice(rank * per, (rank + 1) * per)
    mine = imgs[:, sl].to(dev)
    pmine = tuple(p[:, sl].to(dev) for p in pri)
    res = {}
    for mode in (1, 0, 1):
        capi.check(capi.lib().wm_set_int(m._handle, b"sp_mode", mode))
        for _ in range(2):  # twice: epochs and buffer parity continue across forwards
            outs, _ = m(mine, pmine, cond_flags=[1, 1, 1])
        torch.cuda.synchronize()
        res[mode] = [o.clone() for o in outs]
    enabled, timed_out = m.peer_gather_status()
    assert enabled and not timed_out, (enabled, timed_out)
    ok = True
    for a, b in zip(res[1], res[0]):
        e = ((a - b).norm() / b.norm()).item()
        ok &= e < 2e-3
        print(f"rank {rank}: p2p vs nccl rel fro {e:.3e}", flush=True)
    # single-GPU forward of all views on this rank's GPU
    ref = pkg.B200VisualGeometryTransformer(depth=depth, dino_depth=dino_depth, intermediate_idxs=(0, 2))
    ref.load_state_dict(sd)
    full, _ = ref(imgs.to(dev), tuple(p.to(dev) for p in pri), cond_flags=[1, 1, 1])
    torch.cuda.synchronize()
    for a, f in zip(res[1], full):
        e = ((a - f[:, sl]).norm() / f[:, sl].norm()).item()
        ok &= e < 2e-3
        print(f"rank {rank}: p2p (world {world}) vs single GPU rel fro {e:.3e}", flush=True)
    # prior extraction with the views sharded (priors.extract_priors(group, world, rank)): the pose box is normalised over
    # ALL views of the scene, so the sharded call must return exactly the rows of the whole-scene call
    pose, depth, intr = W.synthetic_scene(per * world, H, Wd, seed=2)
    allv = {"img": imgs.to(dev), "camera_pose": pose.to(dev), "depthmap": depth.to(dev), "camera_intrinsics": intr.to(dev)}
    locv = {k: v[:, sl].contiguous() for k, v in allv.items()}
    pf = pkg.priors.extract_priors(allv)
    pl = pkg.priors.extract_priors(locv, group=None, world=world, rank=rank)
    torch.cuda.synchronize()
    for name, a_, f_ in zip(("depth", "rays", "poses"), pl, pf):
        same = torch.equal(a_, f_[:, sl])
        print(f"rank {rank}: sharded extract_priors {name} == whole-scene rows: {same}", flush=True)
        ok &= same
    # camera head with the views sharded: per-view layers on the local views, q|k|v rows of the trunk all-gathered
    # (B200CameraHead._run) against all views on one GPU; same arithmetic per row, so equal to fp32 rounding at most
    H2 = pkg.heads
    hsd = W.synthetic_heads_state_dict(seed=0)
    cam = H2.B200CameraHead(dim_in=2048)
    cam.load_state_dict({k[len("cam_head."):]: v for k, v in hsd.items() if k.startswith("cam_head.")})
    g = torch.Generator().manual_seed(5)
    ctok = torch.randn(1, per * world, 1, 2048, generator=g).to(dev)
    whole = cam([ctok])
    for _ in range(2):
        mine_c = cam([ctok[:, sl].contiguous()], shard=(None, world, rank))
    torch.cuda.synchronize()
    for a_, f_ in zip(mine_c, whole):
        e = (a_ - f_[:, sl]).abs().max().item()
        ok &= e <= 1e-5
    print(f"rank {rank}: sharded camera head vs all views on one GPU max abs {e:.2e}", flush=True)
    # Gaussian branch with the views sharded: every rank gathers the other ranks' splats and merges the whole scene;
    # the result must equal, bit for bit, the merge of all views done on one GPU (same splats, same order)
    gsd = W.synthetic_heads_state_dict(seed=0, gs=True)
    g = torch.Generator().manual_seed(3)
    S, h, w = per * world, 28, 42
    feat = torch.randn(1, S, h, w, 128, generator=g).to(dev)
    im = torch.rand(1, S, 3, h, w, generator=g).to(dev)
    cp = torch.randn(1, S, 9, generator=g)
    cp[..., 7:] = cp[..., 7:].abs() * 0.3 + 0.6
    preds = {"camera_params": cp.to(dev), "gs_depth": (torch.rand(1, S, h, w, 1, generator=g) * 0.05 + 0.2).to(dev)}
    outs = []
    for sharded, partitioned in ((True, True), (True, False), (False, False)):
        r = H2.B200GaussianSplatRenderer(voxel_size=0.01, sharded=sharded)
        r.partitioned = partitioned
        r.load_state_dict({k[len("gs_renderer."):]: v for k, v in gsd.items() if k.startswith("gs_renderer.")})
        v = sl if sharded else slice(0, S)
        p = {k: t[:, v].contiguous() for k, t in preds.items()}
        res = r.render(feat[:, v].permute(0, 1, 4, 2, 3), im[:, v].contiguous(), p)
        if partitioned:  # this rank merged one key range of the scene: the ranges in rank order are the scene
            assert res.get("splats_partitioned")
            mine_k = res["splats"]["means"][0].shape[0]
            res["splats"] = r.gather_partitions(res["splats"])
        outs.append(res["splats"])
    torch.cuda.synchronize()
    K = outs[2]["means"][0].shape[0]
    same_p = all(torch.equal(outs[0][k][0], outs[2][k][0]) for k in ("means", "sh", "opacities", "scales", "quats"))
    same = all(torch.equal(outs[1][k][0], outs[2][k][0]) for k in ("means", "sh", "opacities", "scales", "quats"))
    print(f"rank {rank}: partitioned merge (this rank: {mine_k} of {K} voxels) == single-GPU merge: {same_p}; "
          f"replicated sharded merge == single-GPU merge: {same} ({S * h * w} splats -> {K} voxels)", flush=True)
    ok &= same and same_p and 0 < K < S * h * w and 0 < mine_k < K
    # the same with the confidence filter: the depth head's confidence is gathered with the splats
    preds["depth_conf"] = (1.0 + torch.randint(0, 50, (1, S, h, w), generator=g).float() * 0.125).to(dev)  # ties on purpose
    outs = []
    for sharded in (True, False):
        r = H2.B200GaussianSplatRenderer(voxel_size=0.01, sharded=sharded, enable_conf_filter=True, conf_threshold_percent=35.0)
        r.load_state_dict({k[len("gs_renderer."):]: v for k, v in gsd.items() if k.startswith("gs_renderer.")})
        v = sl if sharded else slice(0, S)
        p = {k: t[:, v].contiguous() for k, t in preds.items()}
        outs.append(r.render(feat[:, v].permute(0, 1, 4, 2, 3), im[:, v].contiguous(), p)["splats"])
    torch.cuda.synchronize()
    Kf = outs[1]["means"][0].shape[0]
    same = all(torch.equal(outs[0][k][0], outs[1][k][0]) for k in ("means", "sh", "opacities", "scales", "quats"))
    print(f"rank {rank}: sharded filter + merge == single-GPU: {same} ({K} -> {Kf} voxels with the 35 % filter)", flush=True)
    ok &= same and 0 < Kf < K
    # rasterize=True with sharded views: every rank renders ITS views from the whole merged scene
    outs = []
    for sharded in (True, False):
        r = H2.B200GaussianSplatRenderer(voxel_size=0.01, sharded=sharded)
        r.load_state_dict({k[len("gs_renderer."):]: v for k, v in gsd.items() if k.startswith("gs_renderer.")})
        v = sl if sharded else slice(0, S)
        p = {k: t[:, v].contiguous() for k, t in preds.items()}
        outs.append(r.render(feat[:, v].permute(0, 1, 4, 2, 3), im[:, v].contiguous(), p, rasterize=True))
    torch.cuda.synchronize()
    for name in ("rendered_colors", "rendered_depths", "rendered_alphas"):
        a, f = outs[0][name], outs[1][name][:, sl]
        e = (a - f).abs().max().item()
        print(f"rank {rank}: sharded {name} {tuple(a.shape)} vs single GPU max abs {e:.2e}, mean alpha {outs[1]['rendered_alphas'].mean().item():.3f}", flush=True)
        ok &= a.shape == f.shape and e <= 1e-6
    t = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if t.item() != 1:
        raise SystemExit(1)
    if rank == 0:
        print("SP_CHECK_OK", flush=True)


if __name__ == "__main__":
    main()

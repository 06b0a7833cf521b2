"""Run the Gaussian branch once (after a warm-up pass) on synthetic tokens: gs DPT head (fp16 maps), parameter head,
prepare_splats and the voxel merge.  Target of the ncu launch list
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/gs_launches.csv python tools/profile_gs.py
Prints CUDA-event times when run without ncu."""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("hunyuanworld-mirror_b200")
W, H = pkg.weights, pkg.heads

n, size = int(os.environ.get("VIEWS", 8)), 518
hsd = W.synthetic_heads_state_dict(seed=0, gs=True)
sub = lambda p: {k[len(p):]: v for k, v in hsd.items() if k.startswith(p)}
head = H.B200DPTHead(output_dim=2, activation="exp+expp1", is_gsdpt=True)
head.load_state_dict(sub("gs_head."))
rend = H.B200GaussianSplatRenderer()
rend.load_state_dict(sub("gs_renderer."))
g = torch.Generator(device="cpu").manual_seed(0)
P = 7 + (size // 14) ** 2
toks = [(torch.randn(1, n, P, 2048, generator=g) * 0.8).cuda() for _ in range(4)]
imgs = torch.rand(1, n, 3, size, size, generator=g).cuda()
cp = torch.randn(1, n, 9, generator=g)
cp[..., 7:] = cp[..., 7:].abs() * 0.3 + 0.6
cp = cp.cuda()
for it in range(2):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record()
    feat, depth, conf = head(toks, imgs, 7)
    ev[1].record()
    out = rend.render(feat, imgs, {"camera_params": cp, "gs_depth": depth})["splats"]
    ev[2].record()
    torch.cuda.synchronize()
    sp = {k: v[0] for k, v in out.items()}
    c2w, K = rend.prediction_cameras(cp[0], size, size)
    ev2 = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev2[0].record()
    for i in range(0, n, 4):  # chunks of four views, as rasterization.py:223-236
        rend.rasterizer.rasterize_splats(sp["means"], sp["quats"], sp["scales"], sp["opacities"], sp["sh"], c2w[i:i + 4], K[i:i + 4], size, size, sh_degree=0)
    ev2[1].record()
    torch.cuda.synchronize()
    print(f"pass {it}: rasterising {n} views of {size}x{size} from {sp['means'].shape[0]} splats: {ev2[0].elapsed_time(ev2[1]):.3f} ms "
          f"({rend.rasterizer.last['n_isects']} tile intersections in the last chunk)", flush=True)
    print(f"pass {it}: gs DPT head {ev[0].elapsed_time(ev[1]):.3f} ms, parameter head + prepare + merge {ev[1].elapsed_time(ev[2]):.3f} ms, "
          f"{n * size * size} splats -> {out['means'][0].shape[0]} voxels", flush=True)
    # the confidence filter alone (off in WorldMirror's configuration): 30 % dropped, the depth-like confidence of this head
    frend = H.B200GaussianSplatRenderer(enable_conf_filter=True, conf_threshold_percent=30.0)
    N = n * size * size
    raw = {k: torch.rand((N, d) if d > 1 else (N,), generator=g).cuda() for k, d in (("means", 3), ("quats", 4), ("scales", 3), ("opacities", 1), ("sh", 3), ("weights", 1))}
    frend.apply_confidence_filter(raw, conf)
    ev3 = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev3[0].record()
    kept = frend.apply_confidence_filter(raw, conf)
    ev3[1].record()
    torch.cuda.synchronize()
    ms = ev3[0].elapsed_time(ev3[1])
    K = kept["means"].shape[0]
    print(f"pass {it}: confidence filter {N} -> {K} splats: {ms:.3f} ms ({(N * 40 + K * 128) / ms / 1e6:.0f} GB/s of algorithmic traffic)", flush=True)

"""Run the prediction heads once (after a warm-up pass) on synthetic tokens: the target of the ncu launch list
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/heads_launches.csv python tools/profile_heads.py
Prints CUDA-event times per head when run without ncu."""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("hunyuanworld-mirror_b200")
W, H = pkg.weights, pkg.heads

n, size = int(os.environ.get("VIEWS", 8)), 518
hsd = W.synthetic_heads_state_dict(seed=0)
sub = lambda p: {k[len(p):]: v for k, v in hsd.items() if k.startswith(p)}
cam = H.B200CameraHead()
cam.load_state_dict(sub("cam_head."))
dpt = H.B200DPTHead(output_dim=4, activation="inv_log+expp1")
dpt.load_state_dict(sub("pts_head."))
g = torch.Generator(device="cpu").manual_seed(0)
P = 7 + (size // 14) ** 2
toks = [(torch.randn(1, n, P, 2048, generator=g) * 0.8).cuda() for _ in range(4)]
imgs = torch.zeros(1, n, 3, size, size, device="cuda")
for it in range(2):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record()
    cam(toks)
    ev[1].record()
    dpt(toks, imgs, 7)
    ev[2].record()
    torch.cuda.synchronize()
    print(f"pass {it}: camera head {ev[0].elapsed_time(ev[1]):.3f} ms, DPT head {ev[1].elapsed_time(ev[2]):.3f} ms", flush=True)

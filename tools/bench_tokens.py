"""BASELINE.json configs[4]: the aggregator-only microbenchmark - the full alternating frame / global attention stack
(24 + 24 blocks) on synthetic tokens, no DINOv2, no heads.

    python tools/bench_tokens.py --views 8 16 32                      # one GPU
    torchrun --nproc-per-node 8 tools/bench_tokens.py --views 32      # views PER RANK; global attention sequence-parallel

Tokens x ~ N(0, 1) * 0.8 fp32 [N, 1376, 1024], 37 x 37 position grid.  Prints one JSON line per
view count: views/s (whole job), ms per forward, algorithmic TFLOP/s of the stack per GPU (linear 25 165 824 FLOP /
token / block + attention 4 * L_kv * 1024 FLOP / token / block) against the measured sustained bf16 peak."""
import argparse
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, nargs="+", default=[8, 16, 32], help="views per GPU")
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    pkg = importlib.import_module("hunyuanworld-mirror_b200")
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    model = pkg.B200VisualGeometryTransformer()
    model.load_state_dict(pkg.weights.synthetic_state_dict(seed=0))
    if world > 1:
        model.init_sequence_parallel()
    peak = 1400.0
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("bf16_tflops_sustained", peak)
    except Exception:
        pass
    P = 1376
    for n in args.views:
        g = torch.Generator(device="cpu").manual_seed(100 + rank)
        x = (torch.randn(n, P, 1024, generator=g) * 0.8).to(dev)
        for _ in range(2):
            model.forward_tokens(x, 37, 37)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            model.forward_tokens(x, 37, 37)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        T, Tall = n * P, n * world * P
        flops = 24 * T * (2 * 25165824 + 4 * 1024 * (P + Tall))
        if rank == 0:
            tf = flops / (ms.item() * 1e-3) / 1e12
            print(json.dumps({"config": "aggregator-only stack on synthetic tokens (BASELINE configs[4])", "views_per_gpu": n, "n_gpus": world,
                              "views_total": n * world, "ms_per_forward": round(ms.item(), 3), "views_per_s": round(n * world / (ms.item() * 1e-3), 1),
                              "stack_tflops_per_gpu": round(tf, 1), "frac_of_sustained_bf16_peak": round(tf / peak, 3)}), flush=True)
        del x
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

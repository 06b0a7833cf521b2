"""Cost of the device-side attention guard (AttnGuard) at the DINOv2 / frame shape: the bounded-logit kernel alone, the
running-max kernel alone, and the guarded pair of launches when the data fits / does not fit the bound.
    python tools/bench_guard.py [views] [L] [only cases whose name contains this]"""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
capi = importlib.import_module("hunyuanworld-mirror_b200.capi")


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    L = int(sys.argv[2]) if len(sys.argv) > 2 else 1374
    H, T = 16, n * L
    lib, st = capi.lib(), torch.cuda.current_stream().cuda_stream
    g = torch.Generator().manual_seed(0)
    q = (torch.randn(T, 1024, generator=g) * 0.2).to(torch.bfloat16).cuda()
    kv = torch.randn(T, 2048, generator=g).to(torch.bfloat16).cuda()
    o = torch.empty(T, 1024, dtype=torch.bfloat16, device="cuda")
    need = lib.wm_attn_workspace_bytes()
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    small = torch.full((32,), 4.0, dtype=torch.float32, device="cuda").view(torch.int32)
    big = torch.full((32,), 400.0, dtype=torch.float32, device="cuda").view(torch.int32)
    a = (q.data_ptr(), 1024, kv.data_ptr(), kv.data_ptr() + 2048, 2048, o.data_ptr(), 1024, n, L, L, H)
    cases = {
        "bounded (variant 10) alone": lambda: lib.wm_flash_attn_ws(st, *a, 1.0, 10, ws.data_ptr(), need),
        "running-max (variant 4) alone": lambda: lib.wm_flash_attn_ws(st, *a, 0.6931471805599453, 4, ws.data_ptr(), need),
        "guarded (10, 4) one launch, data fits": lambda: lib.wm_flash_attn_guarded(st, *a, 10, 4, ws.data_ptr(), need, small.data_ptr(), 80.0),
        "guarded (10, 4) one launch, does not fit": lambda: lib.wm_flash_attn_guarded(st, *a, 10, 4, ws.data_ptr(), need, big.data_ptr(), 80.0),
        "guarded (12, 7) two launches, data fits": lambda: lib.wm_flash_attn_guarded(st, *a, 12, 7, ws.data_ptr(), need, small.data_ptr(), 80.0),
        "guarded (12, 7) two launches, does not fit": lambda: lib.wm_flash_attn_guarded(st, *a, 12, 7, ws.data_ptr(), need, big.data_ptr(), 80.0),
    }
    flops = 4.0 * n * H * L * L * 64
    only = sys.argv[3] if len(sys.argv) > 3 else ""
    for name, fn in cases.items():
        if only not in name:
            continue
        for _ in range(5):
            capi.check(fn())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(50):
            capi.check(fn())
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 50
        print(f"{name:44s} {us:8.1f} us  {flops / us / 1e6:7.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()

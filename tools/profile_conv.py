"""One 3x3 convolution 256 -> 256 on 8 maps of 148 x 148 (the refinenet1 shape of the DPT heads at 518 px), the target of
    ncu --set full --clock-control none --import-source on -k regex:conv_kernel -s 2 -c 1 -o gpurun_out/r01_conv256_full python tools/profile_conv.py
Without ncu it prints the CUDA-event time and TFLOP/s."""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("hunyuanworld-mirror_b200")
capi, H = pkg.capi, pkg.heads
n, hw, cin, cout = 8, int(os.environ.get("HW", 148)), 256, int(os.environ.get("COUT", 256))
x = torch.randn(n, hw, hw, cin, device="cuda").half()
w = (torch.randn(cout, 9 * cin, device="cuda") * 0.02).half()
b = torch.zeros(cout, device="cuda")
out = torch.empty(n, hw, hw, cout, device="cuda", dtype=torch.float16)
st = torch.cuda.current_stream().cuda_stream
run = lambda: capi.check(capi.lib().wm_conv2d_nhwc(st, x.data_ptr(), n, hw, hw, cin, w.data_ptr(), 9, cout, b.data_ptr(), None, 0, None, 1, out.data_ptr()))
for _ in range(3):
    run()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"conv3x3 {cin}->{cout} on {n} x {hw}x{hw}: {ms:.4f} ms, {2.0 * n * hw * hw * 9 * cin * cout / ms * 1e-9:.1f} TFLOP/s (fp16 operands, fp32 accumulation)")

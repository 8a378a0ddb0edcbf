"""Times gemm_kernel tile widths on the batch-16 conv shapes (CUDA events, L2 flushed between repetitions).
python tools/conv_sweep.py"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
import b200sd  # noqa: E402,F401
from b200sd import ops  # noqa: E402
from b200sd.layers import Kernels  # noqa: E402

dev = torch.device("cuda:0")
K = Kernels(dev)
dt = K.op_dtype
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)
shapes = [(32, 96, 96, 320, 320), (32, 96, 96, 640, 320), (32, 48, 48, 640, 640), (32, 48, 48, 1280, 640),
          (32, 24, 24, 1280, 1280), (32, 12, 12, 1280, 1280), (32, 12, 12, 2560, 1280)]
for n, h, w, cin, cout in shapes:
    a = torch.randn(n, h, w, cin, device=dev).to(dt)
    wt = (torch.randn(cout, 9 * cin, device=dev) / (9 * cin) ** 0.5).to(dt)
    bias = torch.randn(cout, device=dev)
    out = torch.empty(n * h * w, cout, device=dev, dtype=dt)
    flops = 2.0 * n * h * w * cout * 9 * cin
    line = f"conv {h}x{w} {cin}->{cout} M{n*h*w} K{9*cin} |"
    for bn in (0, 64, 128, 160, 256):
        ts = []
        for _ in range(4):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.gemm(a, n, h, w, cin, 9, wt, bias, False, None, 0, None, out, False, bn, K.gemm_ws, 1)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        us = sorted(ts)[1]
        line += f" bn{bn}:{us:7.1f}us {flops / us / 1e6:5.0f}TF |"
    print(line)

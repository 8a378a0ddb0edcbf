"""Times the level-0 3x3 convolution (32 images, 96x96, 320 -> 320: the weakest conv shape of the batch-16 pass) with each
epilogue the UNet gives it, for timing and ncu captures.  python tools/conv_case.py [CASE|all] [reps]
CASE: plain (bias) | conv1 (per-image row bias, column sums) | conv2 (bias + residual, column sums) | conv2_nosums"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
import b200sd  # noqa: E402,F401
from b200sd.layers import Kernels  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda:0")
K = Kernels(dev)
dt = K.op_dtype
n, h, w, cin, cout = 32, 96, 96, 320, 320
a = torch.randn(n, h, w, cin, device=dev).to(dt)
wt = (torch.randn(cout, 9 * cin, device=dev) / (9 * cin) ** 0.5).to(dt)
bias = torch.randn(cout, device=dev)
table = torch.randn(n, 3 * cout, device=dev)
res = torch.randn(n, h, w, cout, device=dev).to(K.trunk_dtype)
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)
CASES = {
    "plain": lambda: K.conv3x3(a, wt, bias, out_dtype=dt),
    "conv1": lambda: K.conv3x3(a, wt, None, row_bias=table[:, cout:2 * cout], out_dtype=dt, gn_next=True),
    "conv2": lambda: K.conv3x3(a, wt, bias, residual=res, gn_next=True),
    "conv2_nosums": lambda: K.conv3x3(a, wt, bias, residual=res),
}
flops = 2.0 * n * h * w * cout * 9 * cin
for name, fn in CASES.items():
    if which not in ("all", name):
        continue
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1000)
    us = sorted(ts)[len(ts) // 2]
    print(f"conv 96x96 320->320 x32 {name:13s}: {us:8.1f} us  {flops / us / 1e6:7.1f} TF/s")

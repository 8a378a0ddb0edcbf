"""Correctness + speed of the CTA-pair GEMM against the single-CTA kernel (B200SD_GEMM_PAIR=0/1 in two processes)."""
import os
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
print("PAIR env:", os.environ.get("B200SD_GEMM_PAIR"))
g = torch.Generator("cpu").manual_seed(0)
# linear cases: (M, N, K)
for m, n, k in [(4096, 1280, 1280), (20000, 640, 640), (18432, 1280, 5120), (5000, 2048, 192)]:
    a = (torch.randn(m, k, generator=g)).to(dev).to(dt)
    w = (torch.randn(n, k, generator=g) * k ** -0.5).to(dev).to(dt)
    b = torch.randn(n, generator=g).to(dev)
    out = K.linear(a, w, b, out_dtype=torch.float32)
    torch.cuda.synchronize()
    ref = a.float() @ w.float().t() + b
    print(f"linear M{m} N{n} K{k}: rel {((out - ref).norm() / ref.norm()).item():.2e}")
# conv cases
for n_img, h, wd, cin, cout in [(4, 24, 24, 1280, 1280), (2, 48, 48, 640, 640), (32, 24, 24, 1280, 1280)]:
    x = torch.randn(n_img, h, wd, cin, generator=g).to(dev).to(dt)
    wt = (torch.randn(cout, cin, 3, 3, generator=g) * (9 * cin) ** -0.5)
    wp = wt.permute(0, 2, 3, 1).reshape(cout, 9 * cin).to(dev).to(dt).contiguous()
    b = torch.randn(cout, generator=g).to(dev)
    out = K.conv3x3(x, wp, b, out_dtype=torch.float32)
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), wt.to(dev).to(dt).float(), b, padding=1).permute(0, 2, 3, 1)
    print(f"conv {n_img}x{h}x{wd} {cin}->{cout}: rel {((out - ref).norm() / ref.norm()).item():.2e}")
    if n_img == 32:
        flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)
        ts = []
        for _ in range(5):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            K.conv3x3(x, wp, b, out_dtype=dt)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        us = sorted(ts)[2]
        print(f"   {us:.1f} us  {2.0 * n_img * h * wd * cout * 9 * cin / us / 1e6:.0f} TFLOP/s")

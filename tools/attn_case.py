"""Times the attention kernel on the UNet's shapes.  python tools/attn_case.py [batch_images] [reps]
(environment switches of the kernel, e.g. B200SD_ATTN_STAGGER=0 or B200SD_ATTN_EXP_EMU=1, apply as usual)"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
import b200sd  # noqa: E402,F401
from b200sd.layers import Kernels  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda:0")
K = Kernels(dev)
dt = K.op_dtype
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)

import os
SHAPES = [(5, 9216, 9216), (10, 2304, 2304), (20, 576, 576), (5, 9216, 77)]
if os.environ.get("B200SD_ATTN_SHAPES") == "big":
    SHAPES = SHAPES[:2]
for heads, sq, skv in SHAPES:
    b = 2 * B
    q = torch.randn(b * sq, heads * 64, device=dev).to(dt)
    k = torch.randn(b * skv, heads * 64, device=dev).to(dt)
    v = torch.randn(b * skv, heads * 64, device=dev).to(dt)
    for _ in range(2):
        K.attention(q, k, v, b, heads, sq, skv, 0.125)
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        K.attention(q, k, v, b, heads, sq, skv, 0.125)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1000)
    ts.sort()
    us = ts[len(ts) // 2]
    flops = 4.0 * b * heads * sq * skv * 64
    print(f"attention b{b} h{heads} sq{sq} skv{skv}: {us:9.1f} us  {flops / us / 1e6:7.1f} TF/s")

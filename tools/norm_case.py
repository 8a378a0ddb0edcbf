"""Times the two norm kernels at the batch-16 level-0 shape (for ncu captures): GroupNorm apply-only (statistics supplied,
as after a producer's column sums) and LayerNorm, 16-bit in and out.  python tools/norm_case.py [gn|ln|all] [reps]"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
import b200sd  # noqa: E402,F401
from b200sd import ops  # noqa: E402
from b200sd.layers import Kernels  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda:0")
K = Kernels(dev)
dt = K.op_dtype
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)


def timed(name, fn, nbytes):
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
    print(f"{name}: {us:8.1f} us  {nbytes / us / 1e3:7.1f} GB/s")


for n, hw, c in [(32, 9216, 320), (32, 2304, 640), (32, 576, 1280), (32, 9216, 640)]:
    x = torch.randn(n, hw, c, device=dev).to(dt)
    gamma, beta = torch.ones(c, device=dev), torch.zeros(c, device=dev)
    if which in ("all", "gn"):
        g = x.float().view(n, hw, 32, c // 32)
        sums = torch.stack([g.sum(dim=(1, 3)), (g * g).sum(dim=(1, 3))], dim=-1).contiguous()
        y = torch.empty_like(x)
        timed(f"groupnorm apply n{n} hw{hw} C{c}", lambda: ops.groupnorm(x, None, n, hw, 32, 1e-5, gamma, beta, True, y, None, None, sums, 0, None),
              2 * x.numel() * 2)
    if which in ("all", "ln"):
        x2 = x.view(n * hw, c)
        y2 = torch.empty_like(x2)
        timed(f"layernorm rows{n * hw} cols{c}", lambda: ops.layernorm(x2, 1e-5, gamma, beta, y2), 2 * x.numel() * 2)

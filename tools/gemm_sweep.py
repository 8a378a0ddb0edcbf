"""Times gemm_kernel variants (block_n x split_k) for the UNet's problem shapes inside a CUDA graph (no CPU launch
overhead, warm L2 -- the in-pipeline condition).  python tools/gemm_sweep.py [B]"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
import b200sd  # noqa: E402
from b200sd import ops  # noqa: E402
from b200sd.layers import Kernels  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dev = torch.device("cuda:0")
K = Kernels(dev)
dt = K.op_dtype


def time_graph(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / (5 * reps)  # us


shapes = [  # (n_img, h, w, cin, taps, n_out, label)
    (1, 1, 1152 * B, 1280, 1, 1280, "linear M1152 K1280"),
    (1, 1, 18432 * B, 320, 1, 320, "linear M18432 K320"),
    (1, 1, 4608 * B, 640, 1, 640, "linear M4608 K640"),
    (1, 1, 288 * B, 1280, 1, 1280, "linear M288 K1280"),
    (1, 1, 1152 * B, 5120, 1, 1280, "linear M1152 K5120"),
    (2 * B, 24, 24, 1280, 9, 1280, "conv 24x24 1280->1280"),
    (2 * B, 12, 12, 1280, 9, 1280, "conv 12x12 1280->1280"),
    (2 * B, 12, 12, 2560, 9, 1280, "conv 12x12 2560->1280"),
    (2 * B, 48, 48, 640, 9, 640, "conv 48x48 640->640"),
    (2 * B, 96, 96, 320, 9, 320, "conv 96x96 320->320"),
]
for n_img, h, w, cin, taps, n_out, label in shapes:
    m = n_img * h * w
    a = torch.randn(n_img, h, w, cin, device=dev).to(dt)
    wt = (torch.randn(n_out, taps * cin, device=dev) / (taps * cin) ** 0.5).to(dt)
    bias = torch.randn(n_out, device=dev)
    res = torch.randn(m, n_out, device=dev)
    out = torch.empty(m, n_out, device=dev)
    flops = 2.0 * m * n_out * taps * cin
    line = f"{label:26s} M{m:6d} N{n_out:5d} K{taps * cin:6d} |"
    best = None
    for bn in (64, 128, 160, 256):
        if n_out % bn:
            continue
        for sk in (1, 2, 3, 4, 6, 8):
            if sk > 1 and (taps * cin // 64) // sk < 2:
                continue
            try:
                us = time_graph(lambda: ops.gemm(a, n_img, h, w, cin, taps, wt, bias, False, None, 0, res, out, False, bn,
                                                 K.gemm_ws, sk))
            except RuntimeError:
                continue
            line += f" bn{bn}/s{sk}:{us:6.1f}"
            if best is None or us < best[0]:
                best = (us, bn, sk)
    auto = time_graph(lambda: ops.gemm(a, n_img, h, w, cin, taps, wt, bias, False, None, 0, res, out, False, 0, K.gemm_ws, 0))
    print(line)
    print(f"{'':26s} -> best {best[0]:.1f} us (bn{best[1]} s{best[2]}) = {flops / best[0] / 1e6:.0f} TFLOP/s; auto {auto:.1f} us")

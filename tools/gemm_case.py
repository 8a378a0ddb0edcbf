"""Runs single GEMM cases from the batch-16 UNet (for timing and for ncu captures).
python tools/gemm_case.py CASE [reps]     CASE in: res320 geglu320 res640 geglu640 qkv320 conv320"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
import b200sd  # noqa: E402,F401
from b200sd.layers import Kernels  # noqa: E402

case = sys.argv[1] if len(sys.argv) > 1 else "res320"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dev = torch.device("cuda:0")
K = Kernels(dev)
dt = K.op_dtype

CASES = {  # name: (M, K, N, residual, geglu, out fp32)
    "res320": (294912, 320, 320, True, False, True),
    "geglu320": (294912, 320, 2560, False, True, False),
    "res640": (73728, 640, 640, True, False, True),
    "geglu640": (73728, 640, 5120, False, True, False),
    "qkv320": (294912, 320, 320, False, False, False),
    "ff2_320": (294912, 1280, 320, True, False, True),
    "ff2_320_out16": (294912, 1280, 320, True, False, False),   # fp32 residual, 16-bit output (the UNet's ff2)
    "res320_inplace": (294912, 320, 320, True, False, True),    # out is the residual: TMA reduce-add
    "res320_16": (294912, 320, 320, True, False, False),        # 16-bit trunk: 16-bit residual and output (staged epilogue)
    "res320_16_inplace": (294912, 320, 320, True, False, False),  # 16-bit trunk, out is the residual: 16-bit TMA reduce-add
    "ff2_320_16_inplace": (294912, 1280, 320, True, False, False),
    "res640_16": (73728, 640, 640, True, False, False),
    "qkv960": (294912, 320, 960, False, False, False),          # fused QKV projection of the 96x96 level
    "res640_16_inplace": (73728, 640, 640, True, False, False),
    "ff2_640_16_inplace": (73728, 2560, 640, True, False, False),
    "qkv1920": (73728, 640, 1920, False, False, False),
}
M, Kd, N, has_res, geglu, out32 = CASES[case]
a = torch.randn(M, Kd, device=dev).to(dt)
w = (torch.randn(N, Kd, device=dev) / Kd ** 0.5).to(dt)
bias = torch.randn(N, device=dev)
res = torch.randn(M, N, device=dev) if has_res else None
if res is not None and "_16" in case:
    res = res.to(dt)
n_cols = N // 2 if geglu else N
out = torch.empty(M, n_cols, device=dev, dtype=torch.float32 if out32 else dt)
if case.endswith("_inplace"):
    out = res
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)


def run():
    K.linear(a, w, bias, residual=res, out=out, geglu=geglu)


for _ in range(3):
    run()
torch.cuda.synchronize()
ts = []
for _ in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1000)
ts.sort()
us = ts[len(ts) // 2]
flops = 2.0 * M * N * Kd
byts = a.numel() * 2 + out.numel() * out.element_size() + (res.numel() * res.element_size() if has_res else 0)
print(f"{case}: M{M} N{N} K{Kd}  {us:8.1f} us  {flops / us / 1e6:7.1f} TF/s  {byts / us / 1e3:7.1f} GB/s (algorithmic)")

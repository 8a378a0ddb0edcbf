"""TEST INFRASTRUCTURE (not collected by pytest): "same box" comparison of SURVEY.md 8(d) -- the reference ships no
Blackwell kernel, its GPU path is stock PyTorch under diffusers (`DEVICE=cuda`, bf16, app/run-sd.py:39,106-116; with
`DEVICE=triton` the UNet is wrapped in torch.compile(fullgraph=True, mode="max-autotune-no-cudagraphs"), :109-116).
diffusers is not installable here, so the stand-in is the oracle's own modules moved to the GPU in bf16, executed by
stock torch (cuDNN / cuBLAS) with torch's fused SDPA kernel in place of the oracle's explicit softmax(QK^T)V (what
diffusers' default attention processor calls), eager or compiled.  Prints one JSON object on the last line.

    python tests/torch_gpu_baseline.py [batch] [--compile] [--size 768] [--steps 25]      # run on the B200 box
bench.py runs it as a subprocess for its `gpu_comparator` block.
"""
import argparse
import json
import sys
import time
from pathlib import Path

import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import sd21 as O  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("batch", nargs="?", type=int, default=16)
ap.add_argument("--compile", action="store_true")
ap.add_argument("--size", type=int, default=768)
ap.add_argument("--steps", type=int, default=25)
ap.add_argument("--explicit", action="store_true", help="also time the oracle's explicit attention (batch <= 4)")
args = ap.parse_args()
B, hw = args.batch, args.size // 8
dev = torch.device("cuda:0")
torch.backends.cudnn.benchmark = True
out = {"batch": B, "dtype": "bf16", "mode": "torch.compile max-autotune-no-cudagraphs" if args.compile else "eager + SDPA",
       "what": f"oracle modules on the GPU under stock torch (cuDNN/cuBLAS/SDPA), {args.size}^2"}


def sdpa_forward(self, x, ctx=None):
    ctx = x if ctx is None else ctx
    b, s, c = x.shape
    d = c // self.heads
    q = self.to_q(x).view(b, s, self.heads, d).transpose(1, 2)
    k = self.to_k(ctx).view(b, ctx.shape[1], self.heads, d).transpose(1, 2)
    v = self.to_v(ctx).view(b, ctx.shape[1], self.heads, d).transpose(1, 2)
    o = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(b, s, c)
    return self.to_out[0](o)


def timed(fn, reps, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


with torch.no_grad(), torch.autocast(device_type="cuda", dtype=torch.bfloat16):
    unet = O.seeded_init_(O.UNet2DCondition(O.sd21_unet_config(sample_size=hw)), 0).eval().to(dev, torch.bfloat16)
    vae = O.seeded_init_(O.VAEDecoder(O.VAEConfig()), 1).eval().to(dev, torch.bfloat16)
    x = torch.randn(2 * B, 4, hw, hw, device=dev, dtype=torch.bfloat16)
    ctx = torch.randn(2 * B, 77, 1024, device=dev, dtype=torch.bfloat16)
    z = torch.randn(min(B, 4), 4, hw, hw, device=dev, dtype=torch.bfloat16)
    t_vae = timed(lambda: vae.decode(z), 2) / z.shape[0]  # ms per image
    out["vae_decode_ms_per_image"] = round(t_vae, 2)
    if args.explicit and B <= 4:  # the explicit S x S scores need 2B * 5 * 9216^2 * 2 bytes
        out["unet_step_ms_explicit_attention"] = round(timed(lambda: unet(x, 961.0, ctx), 3), 2)
    O.Attention.forward = sdpa_forward
    step = unet
    if args.compile:
        unet = unet.to(memory_format=torch.channels_last)  # app/run-sd.py:110
        t0 = time.perf_counter()
        step = torch.compile(unet, fullgraph=True, mode="max-autotune-no-cudagraphs")  # app/run-sd.py:112-116
        step(x, 961.0, ctx)
        torch.cuda.synchronize()
        out["compile_s"] = round(time.perf_counter() - t0, 1)
    t_unet = timed(lambda: step(x, 961.0, ctx), 3)
    out["unet_step_ms"] = round(t_unet, 2)
    per_image = (args.steps * t_unet / B + t_vae)
    out["value"] = out["images_per_s"] = round(1e3 / per_image, 3)
    out["how"] = f"1000 / ({args.steps} x unet_step_ms / batch + vae_decode_ms_per_image)"
print(json.dumps(out))

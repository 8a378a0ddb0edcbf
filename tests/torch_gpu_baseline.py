"""TEST INFRASTRUCTURE (not collected by pytest): "same box" comparison of SURVEY.md 8(d) -- the reference ships no
Blackwell kernel, its GPU path is stock PyTorch under diffusers (`DEVICE=cuda`, bf16, app/run-sd.py:39,106-116).
diffusers is not installable here, so the stand-in is the oracle's own modules moved to the GPU in bf16, executed by
stock torch (cuDNN / cuBLAS), once with the oracle's explicit softmax(QK^T)V and once with torch's fused SDPA kernel
swapped in for it (what diffusers' default attention processor calls).  Prints one JSON object.

    python tests/torch_gpu_baseline.py [batch]      # run on the B200 box
"""
import json
import sys
import time
from pathlib import Path

import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import sd21 as O  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
dev = torch.device("cuda:0")
torch.backends.cudnn.benchmark = True
out = {"batch": B, "dtype": "bf16", "what": "oracle modules on the GPU under stock torch (cuDNN/cuBLAS), 768^2"}


def sdpa_forward(self, x, ctx=None):
    ctx = x if ctx is None else ctx
    b, s, c = x.shape
    d = c // self.heads
    q = self.to_q(x).view(b, s, self.heads, d).transpose(1, 2)
    k = self.to_k(ctx).view(b, ctx.shape[1], self.heads, d).transpose(1, 2)
    v = self.to_v(ctx).view(b, ctx.shape[1], self.heads, d).transpose(1, 2)
    o = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(b, s, c)
    return self.to_out[0](o)


def timed(fn, reps):
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
    unet = O.seeded_init_(O.UNet2DCondition(O.sd21_unet_config(sample_size=96)), 0).eval().to(dev, torch.bfloat16)
    vae = O.seeded_init_(O.VAEDecoder(O.VAEConfig()), 1).eval().to(dev, torch.bfloat16)
    x = torch.randn(2 * B, 4, 96, 96, device=dev, dtype=torch.bfloat16)
    ctx = torch.randn(2 * B, 77, 1024, device=dev, dtype=torch.bfloat16)
    z = torch.randn(min(B, 4), 4, 96, 96, device=dev, dtype=torch.bfloat16)
    t_vae = timed(lambda: vae.decode(z), 2) / z.shape[0]  # ms per image
    out["vae_decode_ms_per_image"] = round(t_vae, 2)
    explicit_ok = B <= 4  # the explicit S x S scores need 2B * 5 * 9216^2 * 2 bytes
    if explicit_ok:
        out["unet_step_ms_explicit_attention"] = round(timed(lambda: unet(x, 961.0, ctx), 3), 2)
    O.Attention.forward = sdpa_forward
    t_sdpa = timed(lambda: unet(x, 961.0, ctx), 3)
    out["unet_step_ms_sdpa"] = round(t_sdpa, 2)
    per_image = (25 * t_sdpa / B + t_vae)
    out["images_per_s_sdpa_25_steps"] = round(1e3 / per_image, 3)
print(json.dumps(out))

"""Generates tests/golden/oracle_tiny.safetensors: SELF-GENERATED regression vectors of the oracle on the tiny
config (the reference ships no golden vectors and diffusers is not installable here -- SURVEY.md 8c).
Run from the repo root:  python tests/golden/make_golden.py"""
import sys
from pathlib import Path

import torch
from safetensors.torch import save_file

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from oracle import sd21 as O  # noqa: E402

unet = O.seeded_init_(O.UNet2DCondition(O.tiny_unet_config()), 0)
vae = O.seeded_init_(O.VAEDecoder(O.tiny_vae_config()), 1)
lat, ctx = O.seeded_inputs(1, 16, 128)
with torch.no_grad():
    eps = unet(torch.cat([lat, lat]), 961.0, ctx)
    final = O.denoise_loop(unet, lat, ctx, 4, 7.5, "ddim")
    img = O.decode_to_image(vae, final)
out = Path(__file__).parent / "oracle_tiny.safetensors"
save_file({"unet_eps_t961": eps.contiguous(), "ddim4_final_latents": final.contiguous(),
           "ddim4_image01": img.contiguous()}, str(out))
print(out, out.stat().st_size, "bytes; eps std", float(eps.std()), "img mean", float(img.mean()))

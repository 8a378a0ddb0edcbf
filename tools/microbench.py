"""SURVEY.md 8(d) config 4 (HBM-bound microbenchmarks at batch 64) and config 1 (512^2, 20 steps, batch 1).
  (i)  K4  CFG + scheduler step on [64, 4, 96, 96], timed over in-graph replays, GB/s against 16 B/element (DDIM)
       and 24 B/element (DPM++2M)
  (ii) VAE decode of [64, 4, 96, 96] (chunks of 4 images), images/s and TFLOP/s
  (iii) config 1 through the pipeline: 512^2, 20 steps, batch 1, DDIM and DPM++2M
python tools/microbench.py   -> one JSON object on stdout"""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
import b200sd  # noqa: E402
from b200sd import ops  # noqa: E402
from b200sd._randinit import random_state_dicts  # noqa: E402

dev = torch.device("cuda:0")
out = {}


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps  # ms


# ---- (i) K4 at batch 64
for kind, bpe in (("ddim", 16), ("dpm++2m", 24)):
    sched = b200sd.B200Scheduler(kind, "v_prediction", dev)
    sched.set_timesteps(25)
    B, n = 64, 64 * 4 * 96 * 96
    x = torch.randn(B, 96, 96, 4, device=dev)
    x0p = torch.zeros_like(x) if sched.needs_x0_prev else None
    eps2 = torch.randn(2 * B, 96, 96, 4, device=dev).to(torch.float16)
    idx = torch.full((1,), 3, device=dev, dtype=torch.int32)
    one = lambda: sched.step_cfg(eps2, idx, x, x0p, 7.5)
    one()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(50):
            one()
    ms = timed(g.replay, 20) / 50
    out[f"k4_{kind}_b64"] = {"us": round(ms * 1e3, 2), "gbs": round(n * bpe / (ms * 1e-3) / 1e9, 1), "bytes_per_elem": bpe,
                             "note": "50 launches per graph replay; the 38-57 MB working set stays in the 126 MB L2 between launches"}

# ---- (ii) VAE decode at batch 64
ucfg = b200sd.UNetConfig(sample_size=96, prediction_type="v_prediction")
vcfg = b200sd.VAEConfig()
usd, vsd = random_state_dicts(ucfg, vcfg, 0)
vae = b200sd.B200VAEDecoder(vsd, vcfg, dev)
z = torch.randn(64, 96, 96, 4, device=dev)
ms = timed(lambda: vae.decode_batched_nhwc(z), 2)
out["vae_decode_b64"] = {"ms": round(ms, 1), "images_per_s": round(64 / (ms * 1e-3), 2),
                         "tflops": round(64 * 5.7543 / (ms * 1e-3), 1)}
del z, usd
torch.cuda.empty_cache()

# ---- (iii) config 1: 512^2, 20 steps, batch 1 (epsilon-prediction base model shapes)
ucfg1 = b200sd.UNetConfig(sample_size=64, prediction_type="epsilon")
usd1, _ = random_state_dicts(ucfg1, vcfg, 0)
unet = b200sd.B200UNet(usd1, ucfg1, dev)
del usd1
for kind in ("ddim", "dpm++2m"):
    pipe = b200sd.B200Pipeline(unet, vae, b200sd.B200Scheduler(kind, "epsilon", dev))
    lat = torch.randn(1, 64, 64, 4, device=dev)
    ctx = torch.randn(2, 77, 1024, device=dev).to(unet.K.op_dtype)
    fn = lambda: pipe.decode(pipe.denoise(lat, ctx, 20), "uint8")
    for _ in range(2):
        fn()
    ms = timed(fn, 5)
    out[f"config1_512_20steps_{kind}"] = {"ms_per_image": round(ms, 2), "images_per_s": round(1e3 / ms, 3),
                                          "tflops": round(34.685 / (ms * 1e-3), 1)}

# ---- (iv) text encoder at the SD2.1 size: 32 prompts (16 conditional + 16 unconditional rows of one batch-16 pass)
cfgc = b200sd.CLIPTextConfig()
g = torch.Generator("cpu").manual_seed(0)
shapes = {"text_model.embeddings.token_embedding.weight": (cfgc.vocab_size, 1024),
          "text_model.embeddings.position_embedding.weight": (77, 1024),
          "text_model.final_layer_norm.weight": (1024,), "text_model.final_layer_norm.bias": (1024,)}
for i in range(cfgc.num_hidden_layers):
    p = f"text_model.encoder.layers.{i}."
    for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
        shapes[p + f"self_attn.{n}.weight"], shapes[p + f"self_attn.{n}.bias"] = (1024, 1024), (1024,)
    for n in ("layer_norm1", "layer_norm2"):
        shapes[p + n + ".weight"], shapes[p + n + ".bias"] = (1024,), (1024,)
    shapes[p + "mlp.fc1.weight"], shapes[p + "mlp.fc1.bias"] = (4096, 1024), (4096,)
    shapes[p + "mlp.fc2.weight"], shapes[p + "mlp.fc2.bias"] = (1024, 4096), (1024,)
sdc = {k: (torch.randn(v, generator=g) * (v[-1] ** -0.5 if len(v) == 2 else 0.1) + (1.0 if k.endswith("norm1.weight") or k.endswith("norm2.weight") or k.endswith("layer_norm.weight") else 0.0))
       for k, v in shapes.items()}
enc = b200sd.B200CLIPTextEncoder(sdc, cfgc, dev)
ids = torch.randint(0, cfgc.vocab_size, (32, 77), generator=g).to(dev)
ms = timed(lambda: enc.encode(ids), 10)
flop = 32 * 77 * cfgc.num_hidden_layers * 2 * (4 * 1024 * 1024 + 2 * 1024 * 4096) + 32 * 16 * cfgc.num_hidden_layers * 4 * 77 * 77 * 64
out["clip_text_b32"] = {"ms": round(ms, 3), "prompts_per_s": round(32 / (ms * 1e-3), 1), "tflops": round(flop / (ms * 1e-3) / 1e12, 1),
                        "note": "eager launches (no CUDA graph): 23 layers x 7 kernels"}
print(json.dumps(out))

"""TEST INFRASTRUCTURE -- CPU fp32 restatement of the SD2.1 text encoder (`pipe.text_encoder`, which the reference
compiles and swaps back in at app/run-sd.py:118-122 and which diffusers calls once per request as
`text_encoder(input_ids, attention_mask=None)[0]`).

The arithmetic lives in the un-vendored dependency `transformers` (`CLIPTextModel`; unpinned in the reference,
app/run-sd.sh:10-12; version 5.5.0 is installed in this image).  PARITY PINNED for this row: `tests/golden/
make_clip_golden.py` runs the real `transformers.CLIPTextModel` in this container on seeded weights / token ids and
commits its output as `tests/golden/clip_tiny.safetensors`; `tests/test_oracle.py` checks this restatement against it.

SD2.1 uses the OpenCLIP ViT-H/14 text tower cut after 23 layers: hidden 1024, 16 heads x 64, MLP 4096 with erf-GELU,
77 learned positions, vocabulary 49408, pre-LayerNorm blocks, causal attention, final LayerNorm; the pipeline takes
`last_hidden_state`.  State-dict keys are transformers' own, so real checkpoints load unchanged.
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict

import torch
import torch.nn.functional as F


@dataclass
class CLIPTextConfig:
    vocab_size: int = 49408
    hidden_size: int = 1024
    intermediate_size: int = 4096
    num_hidden_layers: int = 23
    num_attention_heads: int = 16
    max_position_embeddings: int = 77
    layer_norm_eps: float = 1e-5


def tiny_clip_config() -> CLIPTextConfig:
    return CLIPTextConfig(vocab_size=1000, hidden_size=128, intermediate_size=512, num_hidden_layers=2,
                          num_attention_heads=2)


def param_shapes(cfg: CLIPTextConfig) -> Dict[str, tuple]:
    """transformers' CLIPTextModel state-dict layout (340,387,840 parameters for the SD2.1 configuration)."""
    c, f = cfg.hidden_size, cfg.intermediate_size
    s = {"text_model.embeddings.token_embedding.weight": (cfg.vocab_size, c),
         "text_model.embeddings.position_embedding.weight": (cfg.max_position_embeddings, c)}
    for i in range(cfg.num_hidden_layers):
        p = f"text_model.encoder.layers.{i}."
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            s[p + f"self_attn.{n}.weight"], s[p + f"self_attn.{n}.bias"] = (c, c), (c,)
        for n in ("layer_norm1", "layer_norm2"):
            s[p + n + ".weight"], s[p + n + ".bias"] = (c,), (c,)
        s[p + "mlp.fc1.weight"], s[p + "mlp.fc1.bias"] = (f, c), (f,)
        s[p + "mlp.fc2.weight"], s[p + "mlp.fc2.bias"] = (c, f), (c,)
    s["text_model.final_layer_norm.weight"], s["text_model.final_layer_norm.bias"] = (c,), (c,)
    return s


def seeded_state_dict(cfg: CLIPTextConfig, seed: int = 0) -> Dict[str, torch.Tensor]:
    """weights ~ N(0, 1/fan_in), embeddings ~ N(0, 1), norm weights near 1: activations stay O(1) through 23 layers"""
    g = torch.Generator("cpu").manual_seed(seed)
    sd = {}
    for k, shp in param_shapes(cfg).items():
        if "layer_norm" in k:
            sd[k] = (1.0 + 0.1 * torch.randn(shp, generator=g)) if k.endswith("weight") else 0.1 * torch.randn(shp, generator=g)
        elif "embedding" in k:
            sd[k] = torch.randn(shp, generator=g) * (1.0 if "token" in k else 0.1)
        elif k.endswith("bias"):
            sd[k] = 0.1 * torch.randn(shp, generator=g)
        else:
            sd[k] = torch.randn(shp, generator=g) * shp[1] ** -0.5
    return sd


def clip_text_forward(sd: Dict[str, torch.Tensor], cfg: CLIPTextConfig, input_ids: torch.Tensor) -> torch.Tensor:
    """input_ids int64 [B, L] -> last_hidden_state fp32 [B, L, hidden].  Follows transformers' CLIPTextTransformer:
    token + position embeddings; per layer  h += out_proj(causal_softmax(q k^T / sqrt(d)) v)  on LayerNorm1(h), then
    h += fc2(gelu_erf(fc1(LayerNorm2(h)))); final LayerNorm."""
    B, L = input_ids.shape
    c, nh = cfg.hidden_size, cfg.num_attention_heads
    d = c // nh
    f32 = lambda k: sd[k].float()
    h = f32("text_model.embeddings.token_embedding.weight")[input_ids] + \
        f32("text_model.embeddings.position_embedding.weight")[:L][None]
    mask = torch.full((L, L), float("-inf")).triu(1)
    for i in range(cfg.num_hidden_layers):
        p = f"text_model.encoder.layers.{i}."
        a = F.layer_norm(h, (c,), f32(p + "layer_norm1.weight"), f32(p + "layer_norm1.bias"), cfg.layer_norm_eps)
        q = F.linear(a, f32(p + "self_attn.q_proj.weight"), f32(p + "self_attn.q_proj.bias"))
        k = F.linear(a, f32(p + "self_attn.k_proj.weight"), f32(p + "self_attn.k_proj.bias"))
        v = F.linear(a, f32(p + "self_attn.v_proj.weight"), f32(p + "self_attn.v_proj.bias"))
        q, k, v = (t.view(B, L, nh, d).transpose(1, 2) for t in (q, k, v))
        s = (q @ k.transpose(-1, -2)) * d ** -0.5 + mask
        o = (s.softmax(-1) @ v).transpose(1, 2).reshape(B, L, c)
        h = h + F.linear(o, f32(p + "self_attn.out_proj.weight"), f32(p + "self_attn.out_proj.bias"))
        a = F.layer_norm(h, (c,), f32(p + "layer_norm2.weight"), f32(p + "layer_norm2.bias"), cfg.layer_norm_eps)
        a = F.gelu(F.linear(a, f32(p + "mlp.fc1.weight"), f32(p + "mlp.fc1.bias")))
        h = h + F.linear(a, f32(p + "mlp.fc2.weight"), f32(p + "mlp.fc2.bias"))
    return F.layer_norm(h, (c,), f32("text_model.final_layer_norm.weight"), f32("text_model.final_layer_norm.bias"),
                        cfg.layer_norm_eps)


def seeded_ids(cfg: CLIPTextConfig, batch: int, seed: int = 5) -> torch.Tensor:
    g = torch.Generator("cpu").manual_seed(seed)
    return torch.randint(0, cfg.vocab_size, (batch, cfg.max_position_embeddings), generator=g)

"""Parameter-shape tables of SD2.1's UNet2DConditionModel and AutoencoderKL decoder under diffusers key names
(SURVEY.md A.5), and seeded random-init state_dicts built from them.  There is no network on the serving box, so
benchmarks and the demo host run the real architecture with random weights; with a real checkpoint the same key
tables validate the loaded state_dict."""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch

from .config import UNetConfig, VAEConfig

Shapes = Dict[str, Tuple[int, ...]]


def _conv(s: Shapes, p: str, cin: int, cout: int, k: int):
    s[p + ".weight"], s[p + ".bias"] = (cout, cin, k, k), (cout,)


def _lin(s: Shapes, p: str, cin: int, cout: int, bias: bool = True):
    s[p + ".weight"] = (cout, cin)
    if bias:
        s[p + ".bias"] = (cout,)


def _norm(s: Shapes, p: str, c: int):
    s[p + ".weight"], s[p + ".bias"] = (c,), (c,)


def _resnet(s: Shapes, p: str, cin: int, cout: int, temb: int):
    _norm(s, p + "norm1", cin)
    _conv(s, p + "conv1", cin, cout, 3)
    if temb:
        _lin(s, p + "time_emb_proj", temb, cout)
    _norm(s, p + "norm2", cout)
    _conv(s, p + "conv2", cout, cout, 3)
    if cin != cout:
        _conv(s, p + "conv_shortcut", cin, cout, 1)


def _transformer(s: Shapes, p: str, c: int, ctx: int):
    _norm(s, p + "norm", c)
    _lin(s, p + "proj_in", c, c)
    t = p + "transformer_blocks.0."
    for i in (1, 2, 3):
        _norm(s, t + f"norm{i}", c)
    for a, kv in (("attn1", c), ("attn2", ctx)):
        _lin(s, t + a + ".to_q", c, c, False)
        _lin(s, t + a + ".to_k", kv, c, False)
        _lin(s, t + a + ".to_v", kv, c, False)
        _lin(s, t + a + ".to_out.0", c, c)
    _lin(s, t + "ff.net.0.proj", c, 8 * c)
    _lin(s, t + "ff.net.2", 4 * c, c)
    _lin(s, p + "proj_out", c, c)


def unet_param_shapes(cfg: UNetConfig) -> Shapes:
    s: Shapes = {}
    boc, L, temb, ctx = cfg.block_out_channels, cfg.layers_per_block, cfg.time_embed_dim, cfg.cross_attention_dim
    nb = len(boc)
    _conv(s, "conv_in", cfg.in_channels, boc[0], 3)
    _lin(s, "time_embedding.linear_1", boc[0], temb)
    _lin(s, "time_embedding.linear_2", temb, temb)
    cin, skips = boc[0], [boc[0]]
    for i, co in enumerate(boc):
        for j in range(L):
            _resnet(s, f"down_blocks.{i}.resnets.{j}.", cin if j == 0 else co, co, temb)
            if cfg.down_has_attn[i]:
                _transformer(s, f"down_blocks.{i}.attentions.{j}.", co, ctx)
            skips.append(co)
        if i < nb - 1:
            _conv(s, f"down_blocks.{i}.downsamplers.0.conv", co, co, 3)
            skips.append(co)
        cin = co
    for j in range(2):
        _resnet(s, f"mid_block.resnets.{j}.", boc[-1], boc[-1], temb)
    _transformer(s, "mid_block.attentions.0.", boc[-1], ctx)
    h = boc[-1]
    rev_attn = list(reversed(cfg.down_has_attn))
    for i, co in enumerate(reversed(boc)):
        for j in range(L + 1):
            _resnet(s, f"up_blocks.{i}.resnets.{j}.", h + skips.pop(), co, temb)
            h = co
            if rev_attn[i]:
                _transformer(s, f"up_blocks.{i}.attentions.{j}.", co, ctx)
        if i < nb - 1:
            _conv(s, f"up_blocks.{i}.upsamplers.0.conv", co, co, 3)
    _norm(s, "conv_norm_out", boc[0])
    _conv(s, "conv_out", boc[0], cfg.out_channels, 3)
    return s


def vae_decoder_param_shapes(cfg: VAEConfig) -> Shapes:
    s: Shapes = {}
    rev = list(reversed(cfg.block_out_channels))
    lc = cfg.latent_channels
    _conv(s, "post_quant_conv", lc, lc, 1)
    _conv(s, "decoder.conv_in", lc, rev[0], 3)
    for j in range(2):
        _resnet(s, f"decoder.mid_block.resnets.{j}.", rev[0], rev[0], 0)
    a = "decoder.mid_block.attentions.0."
    _norm(s, a + "group_norm", rev[0])
    for n in ("to_q", "to_k", "to_v", "to_out.0"):
        _lin(s, a + n, rev[0], rev[0])
    cin = rev[0]
    for i, co in enumerate(rev):
        for j in range(cfg.layers_per_block + 1):
            _resnet(s, f"decoder.up_blocks.{i}.resnets.{j}.", cin if j == 0 else co, co, 0)
        if i < len(rev) - 1:
            _conv(s, f"decoder.up_blocks.{i}.upsamplers.0.conv", co, co, 3)
        cin = co
    _norm(s, "decoder.conv_norm_out", rev[-1])
    _conv(s, "decoder.conv_out", rev[-1], cfg.out_channels, 3)
    return s


def _fill(shapes: Shapes, seed: int) -> Dict[str, torch.Tensor]:
    g = torch.Generator("cpu").manual_seed(seed)
    out = {}
    for name in sorted(shapes):
        shp = shapes[name]
        is_norm = "norm" in name.rsplit(".", 2)[-2] if name.count(".") else False
        if name.endswith(".weight") and len(shp) >= 2:
            fan_in = math.prod(shp[1:])
            out[name] = torch.randn(shp, generator=g) / math.sqrt(fan_in)   # variance preserving
        elif name.endswith(".weight"):
            out[name] = 1.0 + 0.1 * torch.randn(shp, generator=g)            # norm gain
        elif is_norm:
            out[name] = 0.1 * torch.randn(shp, generator=g)
        else:
            out[name] = 0.05 * torch.randn(shp, generator=g)
    return out


def random_state_dicts(ucfg: UNetConfig, vcfg: VAEConfig, seed: int = 0):
    return _fill(unet_param_shapes(ucfg), seed), _fill(vae_decoder_param_shapes(vcfg), seed + 1)

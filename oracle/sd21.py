"""fp32 CPU oracle for the SD2.1 denoising hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference`` legs
may import this module.  The product path (``scalable-hw-agnostic-inference_b200/b200sd``) never does.

PARITY UNPINNED.  The reference (``/root/reference``) holds no model math, no tests and no golden
vectors: every FLOP of its hot path runs inside un-vendored, un-pinned ``diffusers``
(``app/run-sd.sh:10`` installs it with ``-U``; call sites ``app/run-sd.py:44,46,107,108,140``), which is
not installed in this image and cannot be installed (no network).  This file therefore restates the
published SD2.1 architecture (``stabilityai/stable-diffusion-2-1`` ``unet/config.json``,
``vae/config.json``, ``scheduler_config.json``; diffusers module semantics as summarised in SURVEY.md
Appendix A) on plain torch fp32 primitives.  What pins the restatement:
  * exact parameter counts (UNet 865,910,724; VAE decoder + post_quant_conv 49,490,199)
  * diffusers ``state_dict`` key names, so real weights load unchanged
  * closed-form scheduler identities (``tests/test_oracle.py``)
  * self-generated golden vectors under ``tests/golden`` (regression pins, NOT reference outputs)

What each piece stands for in the reference's call stack:
  UNet2DCondition    -> ``pipe.unet``          (swapped at app/run-sd.py:110-116)
  VAEDecoder         -> ``pipe.vae.decoder`` + ``pipe.vae.post_quant_conv`` (app/run-sd.py:124-134)
  VAEEncoder         -> ``pipe.vae.encoder`` + ``pipe.vae.quant_conv`` (img2img; SURVEY.md 8f rank 4 -- the reference loads the
                        whole AutoencoderKL at app/run-sd.py:107 but its text2img path never calls encode)
  DDIM               -> ``DDIMScheduler.from_config`` forced at app/run-sd.py:108
  DPMSolverPP2M/Euler-> the schedulers BASELINE.json words ("DPMSolver++/Euler")
  denoise_loop       -> ``StableDiffusionPipeline.__call__`` steps 2-6, call site app/run-sd.py:140
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


# ----------------------------------------------------------------------------------------------
# configs
# ----------------------------------------------------------------------------------------------
@dataclass
class UNetConfig:
    in_channels: int = 4
    out_channels: int = 4
    sample_size: int = 96
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    layers_per_block: int = 2
    # diffusers calls this attention_head_dim but for SD2.1 it holds head COUNTS (head dim is 64)
    attention_head_dim: Tuple[int, ...] = (5, 10, 20, 20)
    cross_attention_dim: int = 1024
    norm_num_groups: int = 32
    norm_eps: float = 1e-5
    # which down blocks carry transformers (CrossAttnDownBlock2D x3, DownBlock2D)
    down_has_attn: Tuple[bool, ...] = (True, True, True, False)
    prediction_type: str = "v_prediction"

    @property
    def time_embed_dim(self) -> int:
        return self.block_out_channels[0] * 4


@dataclass
class VAEConfig:
    latent_channels: int = 4
    out_channels: int = 3
    block_out_channels: Tuple[int, ...] = (128, 256, 512, 512)
    layers_per_block: int = 2
    norm_num_groups: int = 32
    scaling_factor: float = 0.18215


def sd21_unet_config(sample_size: int = 96, prediction_type: str = "v_prediction") -> UNetConfig:
    return UNetConfig(sample_size=sample_size, prediction_type=prediction_type)


def tiny_unet_config() -> UNetConfig:
    """Same topology at toy width: every code path of the real model, seconds on a CPU."""
    return UNetConfig(sample_size=16, block_out_channels=(64, 128, 128, 128), attention_head_dim=(1, 2, 2, 2),
                      cross_attention_dim=128)


def tiny_vae_config() -> VAEConfig:
    return VAEConfig(block_out_channels=(64, 64, 128, 128))


# ----------------------------------------------------------------------------------------------
# UNet2DConditionModel restatement
# ----------------------------------------------------------------------------------------------
def timestep_embedding(t: torch.Tensor, dim: int) -> torch.Tensor:
    """flip_sin_to_cos=True, freq_shift=0: cat([cos, sin])."""
    half = dim // 2
    freqs = torch.exp(-math.log(10000.0) * torch.arange(half, dtype=torch.float32, device=t.device) / half)
    a = t.float()[:, None] * freqs[None, :]
    return torch.cat([torch.cos(a), torch.sin(a)], dim=-1)


class TimestepEmbedding(nn.Module):
    def __init__(self, cin: int, dim: int):
        super().__init__()
        self.linear_1 = nn.Linear(cin, dim)
        self.linear_2 = nn.Linear(dim, dim)

    def forward(self, x):
        return self.linear_2(F.silu(self.linear_1(x)))


class ResnetBlock2D(nn.Module):
    def __init__(self, cin: int, cout: int, temb_dim: Optional[int], groups: int, eps: float):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb_dim, cout) if temb_dim else None
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x, emb=None):
        h = self.conv1(F.silu(self.norm1(x)))
        if self.time_emb_proj is not None:
            h = h + self.time_emb_proj(F.silu(emb))[:, :, None, None]
        h = self.conv2(F.silu(self.norm2(h)))
        skip = self.conv_shortcut(x) if self.conv_shortcut is not None else x
        return skip + h


class Attention(nn.Module):
    """to_q/k/v without bias, to_out.0 with bias, heads of ``dim // heads``; explicit softmax(QK^T/sqrt(d))V."""

    def __init__(self, dim: int, heads: int, ctx_dim: Optional[int] = None, qkv_bias: bool = False):
        super().__init__()
        self.heads = heads
        kv = ctx_dim or dim
        self.to_q = nn.Linear(dim, dim, bias=qkv_bias)
        self.to_k = nn.Linear(kv, dim, bias=qkv_bias)
        self.to_v = nn.Linear(kv, dim, bias=qkv_bias)
        self.to_out = nn.ModuleList([nn.Linear(dim, dim)])

    def forward(self, x, ctx=None):
        ctx = x if ctx is None else ctx
        b, s, c = x.shape
        d = c // self.heads
        q = self.to_q(x).view(b, s, self.heads, d).transpose(1, 2)
        k = self.to_k(ctx).view(b, ctx.shape[1], self.heads, d).transpose(1, 2)
        v = self.to_v(ctx).view(b, ctx.shape[1], self.heads, d).transpose(1, 2)
        p = torch.softmax(q @ k.transpose(-1, -2) * (d ** -0.5), dim=-1)
        o = (p @ v).transpose(1, 2).reshape(b, s, c)
        return self.to_out[0](o)


class GEGLU(nn.Module):
    def __init__(self, dim: int, inner: int):
        super().__init__()
        self.proj = nn.Linear(dim, inner * 2)

    def forward(self, x):
        val, gate = self.proj(x).chunk(2, dim=-1)
        return val * F.gelu(gate)  # erf gelu


class FeedForward(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        # net.0 = GEGLU, net.1 = Dropout (no params), net.2 = Linear -- key names as in diffusers
        self.net = nn.ModuleList([GEGLU(dim, dim * 4), nn.Identity(), nn.Linear(dim * 4, dim)])

    def forward(self, x):
        return self.net[2](self.net[0](x))


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim: int, heads: int, ctx_dim: int):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, heads)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, heads, ctx_dim)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)

    def forward(self, h, ctx):
        h = h + self.attn1(self.norm1(h))
        h = h + self.attn2(self.norm2(h), ctx)
        h = h + self.ff(self.norm3(h))
        return h


class Transformer2DModel(nn.Module):
    """use_linear_projection=True, one layer, GroupNorm eps 1e-6."""

    def __init__(self, dim: int, heads: int, ctx_dim: int, groups: int):
        super().__init__()
        self.norm = nn.GroupNorm(groups, dim, eps=1e-6)
        self.proj_in = nn.Linear(dim, dim)
        self.transformer_blocks = nn.ModuleList([BasicTransformerBlock(dim, heads, ctx_dim)])
        self.proj_out = nn.Linear(dim, dim)

    def forward(self, x, ctx):
        b, c, hh, ww = x.shape
        h = self.norm(x).permute(0, 2, 3, 1).reshape(b, hh * ww, c)
        h = self.proj_in(h)
        h = self.transformer_blocks[0](h, ctx)
        h = self.proj_out(h)
        return h.reshape(b, hh, ww, c).permute(0, 3, 1, 2) + x


class Downsample2D(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, stride=2, padding=1)

    def forward(self, x):
        return self.conv(x)


class Upsample2D(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class DownBlock(nn.Module):
    def __init__(self, cin, cout, temb, layers, groups, eps, heads, ctx_dim, has_attn, add_down):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if j == 0 else cout, cout, temb, groups, eps) for j in range(layers)])
        if has_attn:
            self.attentions = nn.ModuleList([Transformer2DModel(cout, heads, ctx_dim, groups) for _ in range(layers)])
        else:
            self.attentions = None
        self.downsamplers = nn.ModuleList([Downsample2D(cout)]) if add_down else None

    def forward(self, h, emb, ctx):
        outs = []
        for j, r in enumerate(self.resnets):
            h = r(h, emb)
            if self.attentions is not None:
                h = self.attentions[j](h, ctx)
            outs.append(h)
        if self.downsamplers is not None:
            h = self.downsamplers[0](h)
            outs.append(h)
        return h, outs


class MidBlock(nn.Module):
    def __init__(self, c, temb, groups, eps, heads, ctx_dim):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, temb, groups, eps) for _ in range(2)])
        self.attentions = nn.ModuleList([Transformer2DModel(c, heads, ctx_dim, groups)])

    def forward(self, h, emb, ctx):
        h = self.resnets[0](h, emb)
        h = self.attentions[0](h, ctx)
        return self.resnets[1](h, emb)


class UpBlock(nn.Module):
    def __init__(self, cins: Sequence[int], cout, temb, groups, eps, heads, ctx_dim, has_attn, add_up):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(ci, cout, temb, groups, eps) for ci in cins])
        if has_attn:
            self.attentions = nn.ModuleList([Transformer2DModel(cout, heads, ctx_dim, groups) for _ in cins])
        else:
            self.attentions = None
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if add_up else None

    def forward(self, h, skips: List[torch.Tensor], emb, ctx):
        for j, r in enumerate(self.resnets):
            h = r(torch.cat([h, skips.pop()], dim=1), emb)
            if self.attentions is not None:
                h = self.attentions[j](h, ctx)
        if self.upsamplers is not None:
            h = self.upsamplers[0](h)
        return h


class UNet2DCondition(nn.Module):
    def __init__(self, cfg: UNetConfig = UNetConfig()):
        super().__init__()
        self.cfg = cfg
        boc = cfg.block_out_channels
        nb = len(boc)
        temb = cfg.time_embed_dim
        g, eps, L = cfg.norm_num_groups, cfg.norm_eps, cfg.layers_per_block
        self.conv_in = nn.Conv2d(cfg.in_channels, boc[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(boc[0], temb)
        self.down_blocks = nn.ModuleList()
        cin = boc[0]
        skip_ch = [boc[0]]
        for i, co in enumerate(boc):
            last = i == nb - 1
            self.down_blocks.append(DownBlock(cin, co, temb, L, g, eps, cfg.attention_head_dim[i],
                                              cfg.cross_attention_dim, cfg.down_has_attn[i], not last))
            skip_ch += [co] * L + ([] if last else [co])
            cin = co
        self.mid_block = MidBlock(boc[-1], temb, g, eps, cfg.attention_head_dim[-1], cfg.cross_attention_dim)
        self.up_blocks = nn.ModuleList()
        rev = list(reversed(boc))
        rev_heads = list(reversed(cfg.attention_head_dim))
        rev_attn = list(reversed(cfg.down_has_attn))
        h_ch = boc[-1]
        for i, co in enumerate(rev):
            cins = []
            for _ in range(L + 1):
                cins.append(h_ch + skip_ch.pop())
                h_ch = co
            self.up_blocks.append(UpBlock(cins, co, temb, g, eps, rev_heads[i], cfg.cross_attention_dim,
                                          rev_attn[i], i < nb - 1))
        self.conv_norm_out = nn.GroupNorm(g, boc[0], eps=eps)
        self.conv_out = nn.Conv2d(boc[0], cfg.out_channels, 3, padding=1)

    def forward(self, sample: torch.Tensor, timestep, encoder_hidden_states: torch.Tensor) -> torch.Tensor:
        """sample [B2,4,h,w] NCHW, timestep scalar or [B2], ctx [B2,77,D] -> [B2,4,h,w]."""
        b = sample.shape[0]
        t = torch.as_tensor(timestep, device=sample.device).reshape(-1).float()
        if t.numel() == 1:
            t = t.expand(b)
        emb = self.time_embedding(timestep_embedding(t, self.cfg.block_out_channels[0]))
        h = self.conv_in(sample)
        skips = [h]
        for blk in self.down_blocks:
            h, outs = blk(h, emb, encoder_hidden_states)
            skips += outs
        h = self.mid_block(h, emb, encoder_hidden_states)
        for blk in self.up_blocks:
            h = blk(h, skips, emb, encoder_hidden_states)
        return self.conv_out(F.silu(self.conv_norm_out(h)))


# ----------------------------------------------------------------------------------------------
# AutoencoderKL.decode restatement
# ----------------------------------------------------------------------------------------------
class VAEMidAttention(nn.Module):
    def __init__(self, c: int, groups: int):
        super().__init__()
        self.group_norm = nn.GroupNorm(groups, c, eps=1e-6)
        self.to_q = nn.Linear(c, c)
        self.to_k = nn.Linear(c, c)
        self.to_v = nn.Linear(c, c)
        self.to_out = nn.ModuleList([nn.Linear(c, c)])

    def forward(self, x):
        b, c, hh, ww = x.shape
        h = self.group_norm(x).view(b, c, hh * ww).transpose(1, 2)
        q, k, v = self.to_q(h), self.to_k(h), self.to_v(h)
        p = torch.softmax(q @ k.transpose(-1, -2) * (c ** -0.5), dim=-1)
        o = self.to_out[0](p @ v)
        return o.transpose(1, 2).reshape(b, c, hh, ww) + x


class VAEMidBlock(nn.Module):
    def __init__(self, c, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, None, groups, 1e-6) for _ in range(2)])
        self.attentions = nn.ModuleList([VAEMidAttention(c, groups)])

    def forward(self, h):
        return self.resnets[1](self.attentions[0](self.resnets[0](h)))


class VAEUpBlock(nn.Module):
    def __init__(self, cin, cout, layers, groups, add_up):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if j == 0 else cout, cout, None, groups, 1e-6) for j in range(layers)])
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if add_up else None

    def forward(self, h):
        for r in self.resnets:
            h = r(h)
        if self.upsamplers is not None:
            h = self.upsamplers[0](h)
        return h


class Decoder(nn.Module):
    def __init__(self, cfg: VAEConfig):
        super().__init__()
        boc = cfg.block_out_channels
        g = cfg.norm_num_groups
        rev = list(reversed(boc))
        self.conv_in = nn.Conv2d(cfg.latent_channels, rev[0], 3, padding=1)
        self.mid_block = VAEMidBlock(rev[0], g)
        self.up_blocks = nn.ModuleList()
        cin = rev[0]
        for i, co in enumerate(rev):
            self.up_blocks.append(VAEUpBlock(cin, co, cfg.layers_per_block + 1, g, i < len(rev) - 1))
            cin = co
        self.conv_norm_out = nn.GroupNorm(g, boc[0], eps=1e-6)
        self.conv_out = nn.Conv2d(boc[0], cfg.out_channels, 3, padding=1)

    def forward(self, z):
        h = self.mid_block(self.conv_in(z))
        for blk in self.up_blocks:
            h = blk(h)
        return self.conv_out(F.silu(self.conv_norm_out(h)))


class VAEDecoder(nn.Module):
    """``post_quant_conv`` + ``decoder`` under the key names of AutoencoderKL."""

    def __init__(self, cfg: VAEConfig = VAEConfig()):
        super().__init__()
        self.cfg = cfg
        self.post_quant_conv = nn.Conv2d(cfg.latent_channels, cfg.latent_channels, 1)
        self.decoder = Decoder(cfg)

    def decode(self, z):
        return self.decoder(self.post_quant_conv(z))

    forward = decode


# ----------------------------------------------------------------------------------------------
# AutoencoderKL.encode restatement (the VAE encoder of img2img).  diffusers' Encoder:
# conv_in -> 4 DownEncoderBlock2D (2 resnets each, the first three followed by a Downsample2D built with padding=0,
# which pads (0,1,0,1) and then convolves 3x3 stride 2 without padding) -> the same mid block as the decoder ->
# GroupNorm(eps 1e-6) + SiLU -> conv_out to 2*latent channels; AutoencoderKL then applies the 1x1 quant_conv and reads
# the result as (mean, logvar) of a diagonal Gaussian.  Published size: 34,163,592 encoder + 72 quant_conv parameters
# (AutoencoderKL 83,653,863 = encoder + quant_conv + post_quant_conv 20 + decoder 49,490,179).
# ----------------------------------------------------------------------------------------------
class EncoderDownsample2D(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, stride=2, padding=0)

    def forward(self, x):
        return self.conv(F.pad(x, (0, 1, 0, 1), mode="constant", value=0.0))


class VAEDownBlock(nn.Module):
    def __init__(self, cin, cout, layers, groups, add_down):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if j == 0 else cout, cout, None, groups, 1e-6) for j in range(layers)])
        self.downsamplers = nn.ModuleList([EncoderDownsample2D(cout)]) if add_down else None

    def forward(self, h):
        for r in self.resnets:
            h = r(h)
        if self.downsamplers is not None:
            h = self.downsamplers[0](h)
        return h


class Encoder(nn.Module):
    def __init__(self, cfg: VAEConfig):
        super().__init__()
        boc = cfg.block_out_channels
        g = cfg.norm_num_groups
        self.conv_in = nn.Conv2d(cfg.out_channels, boc[0], 3, padding=1)
        self.down_blocks = nn.ModuleList()
        cin = boc[0]
        for i, co in enumerate(boc):
            self.down_blocks.append(VAEDownBlock(cin, co, cfg.layers_per_block, g, i < len(boc) - 1))
            cin = co
        self.mid_block = VAEMidBlock(boc[-1], g)
        self.conv_norm_out = nn.GroupNorm(g, boc[-1], eps=1e-6)
        self.conv_out = nn.Conv2d(boc[-1], 2 * cfg.latent_channels, 3, padding=1)

    def forward(self, x):
        h = self.conv_in(x)
        for blk in self.down_blocks:
            h = blk(h)
        h = self.mid_block(h)
        return self.conv_out(F.silu(self.conv_norm_out(h)))


class VAEEncoder(nn.Module):
    """``encoder`` + ``quant_conv`` under the key names of AutoencoderKL; ``encode`` returns the moments
    [B, 2*latent, h/8, w/8] = cat(mean, logvar) that diffusers wraps in DiagonalGaussianDistribution."""

    def __init__(self, cfg: VAEConfig = VAEConfig()):
        super().__init__()
        self.cfg = cfg
        self.encoder = Encoder(cfg)
        self.quant_conv = nn.Conv2d(2 * cfg.latent_channels, 2 * cfg.latent_channels, 1)

    def encode(self, x):
        return self.quant_conv(self.encoder(x))

    forward = encode


def gaussian_sample(moments: torch.Tensor, noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """DiagonalGaussianDistribution: mean, logvar = chunk(moments, 2, dim=1); logvar clamped to [-30, 20];
    sample = mean + exp(0.5 * logvar) * noise; mode (noise None) = mean."""
    mean, logvar = torch.chunk(moments, 2, dim=1)
    if noise is None:
        return mean
    return mean + torch.exp(0.5 * torch.clamp(logvar, -30.0, 20.0)) * noise


# ----------------------------------------------------------------------------------------------
# deterministic, variance-preserving random init (there are no checkpoints on this box)
# ----------------------------------------------------------------------------------------------
@torch.no_grad()
def seeded_init_(model: nn.Module, seed: int = 0) -> nn.Module:
    """Weights ~ N(0, 1/fan_in), small biases, norm affine near (1, 0): every branch contributes O(1), so
    parity errors are not hidden behind tiny activations."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    for _, m in sorted(model.named_modules(), key=lambda kv: kv[0]):
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            fan_in = m.weight[0].numel()
            m.weight.copy_(torch.randn(m.weight.shape, generator=g) / math.sqrt(fan_in))
            if m.bias is not None:
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.05)
        elif isinstance(m, (nn.GroupNorm, nn.LayerNorm)):
            m.weight.copy_(1.0 + 0.1 * torch.randn(m.weight.shape, generator=g))
            m.bias.copy_(0.1 * torch.randn(m.bias.shape, generator=g))
    return model


# ----------------------------------------------------------------------------------------------
# schedulers.  Every scheduler exposes the per-step update as a LINEAR form
#   x' = cx * x + cm * m + cp * x0_prev ,   x0 = ax * x + am * m     (m = guided model output)
# with host-fp64 scalars -- the same five numbers the CUDA kernel takes -- plus a direct
# (non-linearised) ``step`` used to cross-check the linear form.
# ----------------------------------------------------------------------------------------------
def _alphas_cumprod() -> torch.Tensor:
    betas = torch.linspace(0.00085 ** 0.5, 0.012 ** 0.5, 1000, dtype=torch.float64) ** 2
    return torch.cumprod(1.0 - betas, dim=0)


@dataclass
class StepCoef:
    cx: float      # x'  coefficient on x
    cm: float      # x'  coefficient on guided model output
    cp: float      # x'  coefficient on previous x0 (DPM++ 2M), else 0
    ax: float      # x0  coefficient on x
    am: float      # x0  coefficient on model output
    in_scale: float = 1.0  # scale_model_input for the NEXT UNet call (Euler), else 1


class _SchedulerBase:
    order = 1
    init_noise_sigma = 1.0

    def __init__(self, prediction_type: str = "v_prediction"):
        assert prediction_type in ("v_prediction", "epsilon")
        self.prediction_type = prediction_type
        self.alphas_cumprod = _alphas_cumprod()
        self.timesteps: List[int] = []

    def _x0_coef(self, a: float) -> Tuple[float, float]:
        """x0 = ax*x + am*m for alpha_bar a."""
        sa, sb = math.sqrt(a), math.sqrt(1.0 - a)
        if self.prediction_type == "v_prediction":
            return sa, -sb
        return 1.0 / sa, -sb / sa

    def scale_model_input(self, x, i):
        return x

    def in_scale(self, i: int) -> float:
        return 1.0


class DDIM(_SchedulerBase):
    """eta=0, 'leading' spacing, steps_offset=1, set_alpha_to_one=False, clip_sample=False."""

    def set_timesteps(self, n: int):
        self.n = n
        r = 1000 // n
        self.ratio = r
        self.timesteps = [int(i * r) + 1 for i in range(n)][::-1]

    def coef(self, i: int) -> StepCoef:
        t = self.timesteps[i]
        tp = t - self.ratio
        a = float(self.alphas_cumprod[t])
        ap = float(self.alphas_cumprod[tp]) if tp >= 0 else float(self.alphas_cumprod[0])
        sa, sb, sap, sbp = math.sqrt(a), math.sqrt(1 - a), math.sqrt(ap), math.sqrt(1 - ap)
        ax, am = self._x0_coef(a)
        if self.prediction_type == "v_prediction":
            # x0 = sa x - sb m ; e = sa m + sb x ; x' = sap x0 + sbp e
            cx, cm = sap * sa + sbp * sb, sbp * sa - sap * sb
        else:
            cx, cm = sap / sa, sbp - sap * sb / sa
        return StepCoef(cx, cm, 0.0, ax, am)

    def step(self, m, i, x, x0_prev=None):
        """direct form, for cross-checking ``coef``."""
        t = self.timesteps[i]
        tp = t - self.ratio
        a = self.alphas_cumprod[t]
        ap = self.alphas_cumprod[tp] if tp >= 0 else self.alphas_cumprod[0]
        x, m = x.double(), m.double()
        if self.prediction_type == "v_prediction":
            x0 = a.sqrt() * x - (1 - a).sqrt() * m
            e = a.sqrt() * m + (1 - a).sqrt() * x
        else:
            x0 = (x - (1 - a).sqrt() * m) / a.sqrt()
            e = m
        return (ap.sqrt() * x0 + (1 - ap).sqrt() * e), x0


class DPMSolverPP2M(_SchedulerBase):
    """DPM-Solver++(2M), midpoint, lower_order_final, final sigma 0, 'leading' spacing (pinned here)."""
    order = 1

    def set_timesteps(self, n: int):
        self.n = n
        r = 1000 // n
        self.timesteps = [int(i * r) + 1 for i in range(n)][::-1]
        ac = self.alphas_cumprod
        sig = [math.sqrt((1 - float(ac[t])) / float(ac[t])) for t in self.timesteps] + [0.0]
        self.sigmas = sig

    @staticmethod
    def _alpha_sigma(s: float) -> Tuple[float, float]:
        al = 1.0 / math.sqrt(s * s + 1.0)
        return al, s * al

    def coef(self, i: int) -> StepCoef:
        n = self.n
        s_s, s_t = self.sigmas[i], self.sigmas[i + 1]
        a_s, sg_s = self._alpha_sigma(s_s)
        a_t, sg_t = self._alpha_sigma(s_t)
        ax, am = self._x0_coef(float(self.alphas_cumprod[self.timesteps[i]]))
        last = i == n - 1
        if last:  # final sigma = 0: x' = x0 exactly (h -> inf)
            return StepCoef(ax, am, 0.0, ax, am)
        lam_s, lam_t = math.log(a_s) - math.log(sg_s), math.log(a_t) - math.log(sg_t)
        h = lam_t - lam_s
        k0 = sg_t / sg_s
        k1 = -a_t * math.expm1(-h)
        first_order = i == 0 or (last and n < 15)
        if first_order:
            d0, dp = k1, 0.0
        else:
            a_p, sg_p = self._alpha_sigma(self.sigmas[i - 1])
            lam_p = math.log(a_p) - math.log(sg_p)
            r0 = (lam_s - lam_p) / h
            # x' = k0 x + k1 x0 + 0.5 k1 (x0 - x0p)/r0
            d0, dp = k1 * (1 + 0.5 / r0), -k1 * 0.5 / r0
        return StepCoef(k0 + d0 * ax, d0 * am, dp, ax, am)

    def step(self, m, i, x, x0_prev=None):
        c = self.coef(i)
        x, m = x.double(), m.double()
        x0 = c.ax * x + c.am * m
        s_s, s_t = self.sigmas[i], self.sigmas[i + 1]
        if i == self.n - 1:
            return x0, x0
        a_s, sg_s = self._alpha_sigma(s_s)
        a_t, sg_t = self._alpha_sigma(s_t)
        h = (math.log(a_t) - math.log(sg_t)) - (math.log(a_s) - math.log(sg_s))
        out = (sg_t / sg_s) * x - a_t * math.expm1(-h) * x0
        if i > 0:
            a_p, sg_p = self._alpha_sigma(self.sigmas[i - 1])
            r0 = ((math.log(a_s) - math.log(sg_s)) - (math.log(a_p) - math.log(sg_p))) / h
            d1 = (x0 - x0_prev.double()) / r0
            out = out - 0.5 * a_t * math.expm1(-h) * d1
        return out, x0


class Euler(_SchedulerBase):
    """EulerDiscrete, 'leading' spacing, init_noise_sigma = sqrt(sigma_max^2+1)."""

    def set_timesteps(self, n: int):
        self.n = n
        r = 1000 // n
        self.timesteps = [int(i * r) + 1 for i in range(n)][::-1]
        ac = self.alphas_cumprod
        self.sigmas = [math.sqrt((1 - float(ac[t])) / float(ac[t])) for t in self.timesteps] + [0.0]
        self.init_noise_sigma = math.sqrt(self.sigmas[0] ** 2 + 1.0)

    def in_scale(self, i: int) -> float:
        return 1.0 / math.sqrt(self.sigmas[i] ** 2 + 1.0) if i < self.n else 1.0

    def scale_model_input(self, x, i):
        return x * self.in_scale(i)

    def coef(self, i: int) -> StepCoef:
        s, sn = self.sigmas[i], self.sigmas[i + 1]
        if self.prediction_type == "v_prediction":
            ax, am = 1.0 / (s * s + 1.0), -s / math.sqrt(s * s + 1.0)
        else:
            ax, am = 1.0, -s
        k = (sn - s) / s
        # x' = x + (x - x0) k
        return StepCoef(1.0 + k * (1.0 - ax), -k * am, 0.0, ax, am, self.in_scale(i + 1))

    def step(self, m, i, x, x0_prev=None):
        s, sn = self.sigmas[i], self.sigmas[i + 1]
        x, m = x.double(), m.double()
        if self.prediction_type == "v_prediction":
            x0 = x / (s * s + 1.0) - m * s / math.sqrt(s * s + 1.0)
        else:
            x0 = x - s * m
        return x + (x - x0) / s * (sn - s), x0


SCHEDULERS = {"ddim": DDIM, "dpm++2m": DPMSolverPP2M, "euler": Euler}


def cfg_combine(eps2: torch.Tensor, g: float) -> torch.Tensor:
    u, c = eps2.chunk(2)
    return u + g * (c - u)


def linear_step(x, m, x0_prev, c: StepCoef):
    """the linear form the CUDA kernel implements; fp32 like the kernel."""
    x0 = c.ax * x + c.am * m
    out = c.cx * x + c.cm * m
    if x0_prev is not None and c.cp != 0.0:
        out = out + c.cp * x0_prev
    return out, x0


# ----------------------------------------------------------------------------------------------
# pipeline loop on given latents + text embeddings
# ----------------------------------------------------------------------------------------------
@torch.no_grad()
def denoise_loop(unet: UNet2DCondition, latents: torch.Tensor, ctx2: torch.Tensor, steps: int,
                 guidance: float = 7.5, scheduler: str = "ddim", return_eps: bool = False):
    """latents [B,4,h,w] unit-variance noise; ctx2 = cat([uncond, cond]) [2B,77,D].  Returns final latents."""
    sch = SCHEDULERS[scheduler](unet.cfg.prediction_type)
    sch.set_timesteps(steps)
    x = latents.float() * sch.init_noise_sigma
    x0_prev = None
    eps_trace = []
    for i, t in enumerate(sch.timesteps):
        x2 = sch.scale_model_input(torch.cat([x, x]), i)
        eps2 = unet(x2, float(t), ctx2)
        m = cfg_combine(eps2, guidance)
        if return_eps:
            eps_trace.append(m)
        x, x0_prev = linear_step(x, m, x0_prev, sch.coef(i))
    return (x, eps_trace) if return_eps else x


@torch.no_grad()
def decode_to_image(vae: VAEDecoder, latents: torch.Tensor) -> torch.Tensor:
    """vae.decode(latents / scaling_factor) -> (img/2+0.5).clamp(0,1) NHWC fp32 in [0,1]."""
    img = vae.decode(latents / vae.cfg.scaling_factor)
    return (img / 2 + 0.5).clamp(0, 1).permute(0, 2, 3, 1).contiguous()


def to_uint8(img01: torch.Tensor) -> torch.Tensor:
    """numpy_to_pil rule: (x*255).round().astype(uint8)."""
    return (img01 * 255.0).round().to(torch.uint8)


@torch.no_grad()
def text2img(unet, vae, latents, ctx2, steps, guidance=7.5, scheduler="ddim"):
    return decode_to_image(vae, denoise_loop(unet, latents, ctx2, steps, guidance, scheduler))


def seeded_inputs(batch: int, hw: int, ctx_dim: int = 1024, seed: int = 1234, ctx_len: int = 77):
    """identical on every machine: latents [B,4,hw,hw], ctx2 [2B,ctx_len,ctx_dim] (uncond first)."""
    lat = torch.stack([torch.randn(4, hw, hw, generator=torch.Generator("cpu").manual_seed(seed + i))
                       for i in range(batch)])
    g = torch.Generator("cpu").manual_seed(seed + 100003)
    unc = torch.randn(1, ctx_len, ctx_dim, generator=g).expand(batch, -1, -1)
    cond = torch.randn(batch, ctx_len, ctx_dim, generator=g)
    return lat, torch.cat([unc, cond]).contiguous()


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def psnr(a01: torch.Tensor, b01: torch.Tensor) -> float:
    mse = float(((a01.double() - b01.double()) ** 2).mean())
    return 99.0 if mse == 0 else 10.0 * math.log10(1.0 / mse)

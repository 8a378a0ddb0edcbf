"""Model configs (the fields of SD2.1's unet/config.json and vae/config.json that the hot path reads)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple


@dataclass
class UNetConfig:
    in_channels: int = 4
    out_channels: int = 4
    sample_size: int = 96
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    layers_per_block: int = 2
    attention_head_dim: Tuple[int, ...] = (5, 10, 20, 20)  # head COUNTS for SD2.1 (head dim = 64)
    cross_attention_dim: int = 1024
    norm_num_groups: int = 32
    norm_eps: float = 1e-5
    down_has_attn: Tuple[bool, ...] = (True, True, True, False)
    prediction_type: str = "v_prediction"
    time_cond_proj_dim = None

    @property
    def time_embed_dim(self) -> int:
        return self.block_out_channels[0] * 4

    @classmethod
    def from_any(cls, cfg) -> "UNetConfig":
        """accepts this class, the oracle's dataclass, or a diffusers-style config object / dict."""
        if isinstance(cfg, cls):
            return cfg
        get = (lambda k, d=None: cfg.get(k, d)) if isinstance(cfg, dict) else (lambda k, d=None: getattr(cfg, k, d))
        out = cls()
        for f in ("in_channels", "out_channels", "sample_size", "layers_per_block", "cross_attention_dim",
                  "norm_num_groups", "norm_eps", "prediction_type"):
            v = get(f)
            if v is not None:
                setattr(out, f, v)
        for f in ("block_out_channels", "attention_head_dim", "down_has_attn"):
            v = get(f)
            if v is not None:
                setattr(out, f, tuple(v))
        dbt = get("down_block_types")
        if dbt is not None:
            out.down_has_attn = tuple("CrossAttn" in t for t in dbt)
        return out


@dataclass
class VAEConfig:
    latent_channels: int = 4
    out_channels: int = 3
    block_out_channels: Tuple[int, ...] = (128, 256, 512, 512)
    layers_per_block: int = 2
    norm_num_groups: int = 32
    scaling_factor: float = 0.18215

    @classmethod
    def from_any(cls, cfg) -> "VAEConfig":
        if isinstance(cfg, cls):
            return cfg
        get = (lambda k, d=None: cfg.get(k, d)) if isinstance(cfg, dict) else (lambda k, d=None: getattr(cfg, k, d))
        out = cls()
        for f in ("latent_channels", "out_channels", "layers_per_block", "norm_num_groups", "scaling_factor"):
            v = get(f)
            if v is not None:
                setattr(out, f, v)
        v = get("block_out_channels")
        if v is not None:
            out.block_out_channels = tuple(v)
        return out

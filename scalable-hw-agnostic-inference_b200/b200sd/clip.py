"""B200CLIPTextEncoder -- drop-in for `pipe.text_encoder`, which the reference compiles and swaps back into the
pipeline at app/run-sd.py:118-122 and which diffusers calls once per request as
`text_encoder(input_ids, attention_mask=None)[0]`.

SD2.1's text encoder is transformers' `CLIPTextModel` with the OpenCLIP ViT-H/14 text tower cut after 23 layers:
pre-LayerNorm blocks, causal attention (16 heads x 64), erf-GELU MLP, final LayerNorm, `last_hidden_state` out.
Every op runs on the hot path's kernels: token + position embedding (`embed_tokens`), LayerNorm, fused QKV GEMM,
flash attention with the causal flag, out-proj GEMM with the residual in the epilogue, fc1 GEMM with the GELU in the
epilogue, fc2 GEMM with the residual.  Weights come in transformers' state-dict layout, so a real checkpoint's
`text_encoder` loads unchanged.  fp32 residual trunk, 16-bit tensor-core operands (as in the UNet).
"""
from __future__ import annotations

from types import SimpleNamespace
from typing import Dict, Optional

import torch

from . import ops
from .layers import Kernels


class CLIPTextConfig(SimpleNamespace):
    def __init__(self, vocab_size=49408, hidden_size=1024, intermediate_size=4096, num_hidden_layers=23,
                 num_attention_heads=16, max_position_embeddings=77, layer_norm_eps=1e-5, **extra):
        super().__init__(vocab_size=vocab_size, hidden_size=hidden_size, intermediate_size=intermediate_size,
                         num_hidden_layers=num_hidden_layers, num_attention_heads=num_attention_heads,
                         max_position_embeddings=max_position_embeddings, layer_norm_eps=layer_norm_eps, **extra)

    @classmethod
    def from_any(cls, c):
        if isinstance(c, cls):
            return c
        keys = ("vocab_size", "hidden_size", "intermediate_size", "num_hidden_layers", "num_attention_heads",
                "max_position_embeddings", "layer_norm_eps")
        get = (lambda k: c[k]) if isinstance(c, dict) else (lambda k: getattr(c, k))
        return cls(**{k: get(k) for k in keys})


class CLIPTextOutput(tuple):
    """what `text_encoder(...)` returns: indexable like transformers' ModelOutput ([0] = last_hidden_state)"""

    def __new__(cls, last_hidden_state):
        self = super().__new__(cls, (last_hidden_state,))
        self.last_hidden_state = last_hidden_state
        return self


class B200CLIPTextEncoder(torch.nn.Module):
    def __init__(self, state_dict: Dict[str, torch.Tensor], config=None, device="cuda", op_dtype=None,
                 out_dtype: Optional[torch.dtype] = None):
        super().__init__()
        self.config = CLIPTextConfig.from_any(config if config is not None else CLIPTextConfig())
        cfg = self.config
        self._device = torch.device(device)
        if self._device.type != "cuda":
            raise RuntimeError("B200CLIPTextEncoder needs a CUDA device (there is no CPU path)")
        ops.check_device(self._device.index or 0)
        if cfg.hidden_size % cfg.num_attention_heads or cfg.hidden_size // cfg.num_attention_heads != 64:
            raise ValueError("head_dim must be 64 (hidden_size / num_attention_heads)")
        self.K = Kernels(self._device, torch.float32, op_dtype)
        OP = self.K.op_dtype
        self.out_dtype = out_dtype or OP
        dev, sd = self._device, state_dict
        f = lambda k: sd[k].to(dev).float().contiguous()
        h = lambda k: sd[k].to(dev).to(OP).contiguous()
        self.tok = f("text_model.embeddings.token_embedding.weight")
        self.pos = f("text_model.embeddings.position_embedding.weight")
        self.layers = []
        for i in range(cfg.num_hidden_layers):
            p = f"text_model.encoder.layers.{i}."
            qkv_w = torch.cat([sd[p + f"self_attn.{n}_proj.weight"] for n in "qkv"]).to(dev).to(OP).contiguous()
            qkv_b = torch.cat([sd[p + f"self_attn.{n}_proj.bias"] for n in "qkv"]).to(dev).float().contiguous()
            self.layers.append(SimpleNamespace(
                ln1=(f(p + "layer_norm1.weight"), f(p + "layer_norm1.bias")),
                ln2=(f(p + "layer_norm2.weight"), f(p + "layer_norm2.bias")),
                qkv_w=qkv_w, qkv_b=qkv_b,
                out_w=h(p + "self_attn.out_proj.weight"), out_b=f(p + "self_attn.out_proj.bias"),
                fc1_w=h(p + "mlp.fc1.weight"), fc1_b=f(p + "mlp.fc1.bias"),
                fc2_w=h(p + "mlp.fc2.weight"), fc2_b=f(p + "mlp.fc2.bias")))
        self.final_ln = (f("text_model.final_layer_norm.weight"), f("text_model.final_layer_norm.bias"))

    @property
    def dtype(self):
        return self.out_dtype

    @property
    def device(self):
        return self._device

    @torch.no_grad()
    def encode(self, input_ids: torch.Tensor) -> torch.Tensor:
        """input_ids integer [B, L] (L <= max_position_embeddings) -> last_hidden_state [B, L, hidden], `out_dtype`."""
        cfg, K = self.config, self.K
        B, L = input_ids.shape
        if L > cfg.max_position_embeddings:
            raise ValueError(f"sequence length {L} > {cfg.max_position_embeddings}")
        ids = input_ids.to(self._device)
        if int(ids.min()) < 0 or int(ids.max()) >= cfg.vocab_size:
            raise ValueError("token id outside the vocabulary")
        ids = ids.to(torch.int32).contiguous()
        c, nh, eps = cfg.hidden_size, cfg.num_attention_heads, cfg.layer_norm_eps
        hs = torch.empty(B * L, c, device=self._device, dtype=torch.float32)
        ops.embed_tokens(ids, self.tok, self.pos, hs)
        for ly in self.layers:
            a = K.layernorm(hs, ly.ln1[0], ly.ln1[1], eps)
            qkv = K.linear(a, ly.qkv_w, ly.qkv_b, out_dtype=K.op_dtype)  # [B*L, 3c]
            o = K.attention(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], B, nh, L, L, 0.125, causal=True)
            hs = K.linear(o, ly.out_w, ly.out_b, residual=hs, out=hs)
            a = K.layernorm(hs, ly.ln2[0], ly.ln2[1], eps)
            a = K.linear(a, ly.fc1_w, ly.fc1_b, out_dtype=K.op_dtype, gelu=True)
            hs = K.linear(a, ly.fc2_w, ly.fc2_b, residual=hs, out=hs)
        # the LayerNorm kernel emits the 16-bit operand format (what the UNet's cross-attention K/V GEMM consumes);
        # a caller that asked for another dtype gets that result converted
        out = K.layernorm(hs, self.final_ln[0], self.final_ln[1], eps)
        if out.dtype != self.out_dtype:
            out = out.to(self.out_dtype)
        return out.view(B, L, c)

    def forward(self, input_ids: torch.Tensor, attention_mask=None, **_kw) -> CLIPTextOutput:
        """`pipe.text_encoder(input_ids, attention_mask=None)`; SD2.1's pipeline passes no mask (the tower is causal and
        the pipeline keeps every position, padding included)."""
        return CLIPTextOutput(self.encode(input_ids))


class HashTokenizer:
    """PLACEHOLDER for transformers' CLIPTokenizer: the CLIP BPE vocabulary is a checkpoint asset that does not exist on
    this box (no network).  Words are lower-cased, split on non-alphanumerics and mapped to a stable id in
    [1, vocab - 3] by SHA-256; framing follows the SD2.1 tokenizer (BOS 49406, EOS 49407, pad id 0, length 77,
    truncation).  With a real checkpoint pass `CLIPTokenizer.from_pretrained(...)`-style callables instead."""

    def __init__(self, vocab_size: int = 49408, length: int = 77):
        self.vocab_size, self.length = vocab_size, length
        self.bos, self.eos, self.pad = vocab_size - 2, vocab_size - 1, 0

    def __call__(self, prompts) -> torch.Tensor:
        import hashlib
        import re
        rows = []
        for p in prompts:
            words = [w for w in re.split(r"[^0-9a-z]+", p.lower()) if w]
            ids = [1 + int.from_bytes(hashlib.sha256(w.encode()).digest()[:8], "little") % (self.vocab_size - 3)
                   for w in words][: self.length - 2]
            row = [self.bos] + ids + [self.eos]
            rows.append(row + [self.pad] * (self.length - len(row)))
        return torch.tensor(rows, dtype=torch.int64)


class PromptEncoder:
    """prompts -> [B, 77, hidden] on the device: tokenizer + B200CLIPTextEncoder, with the per-prompt results cached
    (the empty negative prompt is encoded once)."""

    def __init__(self, encoder: B200CLIPTextEncoder, tokenizer=None, cache_size: int = 256):
        self.encoder = encoder
        self.tokenizer = tokenizer or HashTokenizer(encoder.config.vocab_size, encoder.config.max_position_embeddings)
        self._cache: Dict[str, torch.Tensor] = {}
        self._cache_size = cache_size

    def _ids(self, prompts):
        t = self.tokenizer(prompts)
        if not torch.is_tensor(t):  # a transformers tokenizer returns a BatchEncoding
            t = self.tokenizer(prompts, padding="max_length", truncation=True,
                               max_length=self.encoder.config.max_position_embeddings, return_tensors="pt").input_ids
        return t

    def __call__(self, prompts) -> torch.Tensor:
        prompts = list(prompts)
        # this call's embeddings are collected locally, so evicting the cache can never lose one of them
        got = {p: self._cache[p] for p in dict.fromkeys(prompts) if p in self._cache}
        todo = [p for p in dict.fromkeys(prompts) if p not in got]
        if todo:
            emb = self.encoder.encode(self._ids(todo))
            for p, e in zip(todo, emb):
                got[p] = e
            if len(self._cache) + len(todo) > self._cache_size:
                self._cache = {p: e for p, e in got.items() if p in self._cache}  # keep what this call still uses
            for p in todo[: max(0, self._cache_size - len(self._cache))]:
                self._cache[p] = got[p]
        return torch.stack([got[p] for p in prompts])

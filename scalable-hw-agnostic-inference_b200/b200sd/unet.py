"""B200UNet -- drop-in for `pipe.unet` (diffusers UNet2DConditionModel; swapped at app/run-sd.py:110-116).

Same call signature and config attributes the pipeline reads; internally NHWC, 16-bit tensor-core operands (fp16 by
default, bf16 selectable) and a 16-bit residual trunk, every FLOP in libb200sd's sm_100a kernels (implicit-GEMM conv /
linear on tcgen05 whose epilogues also emit the next GroupNorm's statistics, flash attention, GroupNorm / LayerNorm).
Weights come from a diffusers-layout state_dict and are packed once at construction.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import ops
from .config import UNetConfig
from .layers import Kernels, ResnetParams, keep_tile_sums, pack_conv3x3, pack_conv3x3_im2col, pack_geglu, run_resnet


class _TransformerParams:
    def __init__(self, sd, prefix: str, dev, dim: int, heads: int, BF: torch.dtype):
        g = lambda k: sd[prefix + k].to(dev)
        f = lambda k: g(k).float().contiguous()
        b = lambda k: g(k).to(BF).contiguous()
        self.dim, self.heads = dim, heads
        self.gn_g, self.gn_b = f("norm.weight"), f("norm.bias")
        self.pin_w, self.pin_b = b("proj_in.weight"), f("proj_in.bias")
        self.pout_w, self.pout_b = b("proj_out.weight"), f("proj_out.bias")
        t = "transformer_blocks.0."
        self.ln = [(f(t + f"norm{i}.weight"), f(t + f"norm{i}.bias")) for i in (1, 2, 3)]
        self.qkv_w = torch.cat([g(t + "attn1.to_q.weight"), g(t + "attn1.to_k.weight"),
                                g(t + "attn1.to_v.weight")]).to(BF).contiguous()
        self.o1_w, self.o1_b = b(t + "attn1.to_out.0.weight"), f(t + "attn1.to_out.0.bias")
        self.q2_w = b(t + "attn2.to_q.weight")
        self.kv2_w = torch.cat([g(t + "attn2.to_k.weight"), g(t + "attn2.to_v.weight")]).to(BF).contiguous()
        self.o2_w, self.o2_b = b(t + "attn2.to_out.0.weight"), f(t + "attn2.to_out.0.bias")
        w1, b1 = g(t + "ff.net.0.proj.weight"), g(t + "ff.net.0.proj.bias")
        self.ff_block_n = ops.pick_block_n(w1.shape[0], True)
        self.ff1_w, self.ff1_b = pack_geglu(w1, b1, self.ff_block_n, BF)
        self.ff2_w, self.ff2_b = b(t + "ff.net.2.weight"), f(t + "ff.net.2.bias")


class B200UNet(torch.nn.Module):
    def __init__(self, state_dict: Dict[str, torch.Tensor], config=None, device="cuda",
                 trunk_dtype: Optional[torch.dtype] = None, op_dtype: Optional[torch.dtype] = None):
        super().__init__()
        self.config = UNetConfig.from_any(config if config is not None else UNetConfig())
        cfg = self.config
        self._device = torch.device(device)
        if self._device.type != "cuda":
            raise RuntimeError("B200UNet needs a CUDA device (there is no CPU path)")
        ops.check_device(self._device.index or 0)
        self.K = Kernels(self._device, trunk_dtype, op_dtype)
        BF = self.K.op_dtype
        self._dtype = BF
        dev, sd = self._device, state_dict
        boc, L = cfg.block_out_channels, cfg.layers_per_block
        nb = len(boc)
        f = lambda k: sd[k].to(dev).float().contiguous()

        # ---- time embedding: one table [rows, sum Cout] per step feeds every resnet's conv1 epilogue
        self.te_w1, self.te_b1 = sd["time_embedding.linear_1.weight"].to(dev).to(BF).contiguous(), f("time_embedding.linear_1.bias")
        self.te_w2, self.te_b2 = sd["time_embedding.linear_2.weight"].to(dev).to(BF).contiguous(), f("time_embedding.linear_2.bias")
        self._temb_w: List[torch.Tensor] = []
        self._temb_b: List[torch.Tensor] = []
        self._temb_off = 0

        def resnet(prefix: str) -> ResnetParams:
            w = sd[prefix + "time_emb_proj.weight"].to(dev)
            bias = sd[prefix + "time_emb_proj.bias"].to(dev).float() + sd[prefix + "conv1.bias"].to(dev).float()
            lo = self._temb_off
            self._temb_off += w.shape[0]
            self._temb_w.append(w.to(BF))
            self._temb_b.append(bias)
            return ResnetParams(sd, prefix, dev, BF, (lo, self._temb_off))

        self.conv_in_w = pack_conv3x3_im2col(sd["conv_in.weight"].to(dev), 64, BF)
        self.conv_in_b = f("conv_in.bias")
        self.down = []
        for i, co in enumerate(boc):
            blk = {"res": [resnet(f"down_blocks.{i}.resnets.{j}.") for j in range(L)], "attn": None, "down": None}
            if cfg.down_has_attn[i]:
                blk["attn"] = [_TransformerParams(sd, f"down_blocks.{i}.attentions.{j}.", dev, co,
                                                  cfg.attention_head_dim[i], BF) for j in range(L)]
            if i < nb - 1:
                p = f"down_blocks.{i}.downsamplers.0.conv."
                blk["down"] = (pack_conv3x3(sd[p + "weight"].to(dev), BF), f(p + "bias"))
            self.down.append(blk)
        self.mid_res = [resnet(f"mid_block.resnets.{j}.") for j in range(2)]
        self.mid_attn = _TransformerParams(sd, "mid_block.attentions.0.", dev, boc[-1], cfg.attention_head_dim[-1], BF)
        self.up = []
        rev = list(reversed(boc))
        rev_heads = list(reversed(cfg.attention_head_dim))
        rev_attn = list(reversed(cfg.down_has_attn))
        for i, co in enumerate(rev):
            blk = {"res": [resnet(f"up_blocks.{i}.resnets.{j}.") for j in range(L + 1)], "attn": None, "up": None}
            if rev_attn[i]:
                blk["attn"] = [_TransformerParams(sd, f"up_blocks.{i}.attentions.{j}.", dev, co, rev_heads[i], BF)
                               for j in range(L + 1)]
            if i < nb - 1:
                p = f"up_blocks.{i}.upsamplers.0.conv."
                blk["up"] = (pack_conv3x3(sd[p + "weight"].to(dev), BF), f(p + "bias"))
            self.up.append(blk)
        self.out_g, self.out_b = f("conv_norm_out.weight"), f("conv_norm_out.bias")
        self.conv_out_w = pack_conv3x3(sd["conv_out.weight"].to(dev), BF)
        self.conv_out_b = f("conv_out.bias")
        self.temb_all_w = torch.cat(self._temb_w).contiguous()
        self.temb_all_b = torch.cat(self._temb_b).contiguous()
        del self._temb_w, self._temb_b
        self._ctx_cache = None
        # proj_out adds onto the block's input in place.  With its column sums (default) it takes the staged epilogue, which
        # reads the residual; without them (B200SD_POUT_SUMS=0) a 16-bit trunk takes the TMA reduce-add and the next
        # GroupNorm runs its own statistics pass.
        import os
        self._pout_sums = os.environ.get("B200SD_POUT_SUMS", "1") != "0"

    # ---- attributes the diffusers pipeline reads -------------------------------------------------------
    @property
    def dtype(self):
        return self._dtype

    @property
    def device(self):
        return self._device

    def set_attention_slice(self, *_a, **_k):
        """pipe.enable_attention_slicing() (app/run-sd.py:135) lands here: flash attention never
        materialises the score matrix, so there is nothing to slice."""
        return None

    def transformers(self) -> List[_TransformerParams]:
        out = []
        for blk in self.down:
            out += blk["attn"] or []
        out.append(self.mid_attn)
        for blk in self.up:
            out += blk["attn"] or []
        return out

    # ---- pieces ---------------------------------------------------------------------------------------
    def project_context(self, ctx: torch.Tensor) -> List[torch.Tensor]:
        """cross-attention K/V of every transformer for a prompt batch; step-invariant, so computed once."""
        BF = self.K.op_dtype
        ctx_bf = ctx.to(device=self._device, dtype=BF).contiguous()
        n, s, d = ctx_bf.shape
        flat = ctx_bf.view(n * s, d)
        return [self.K.linear(flat, T.kv2_w, None, out_dtype=BF) for T in self.transformers()]

    def time_table(self, t: torch.Tensor) -> torch.Tensor:
        """t: device fp32 [rows] -> fp32 [rows, sum Cout] = time_emb_proj(silu(emb)) + conv1.bias for all resnets."""
        rows = t.numel()
        c0 = self.config.block_out_channels[0]
        d = self.config.time_embed_dim
        h1 = torch.empty(rows, d, device=self._device, dtype=torch.float32)
        ops.small_linear(t, rows, c0, c0, False, True, self.te_w1, self.te_b1, h1)
        emb = torch.empty_like(h1)
        ops.small_linear(h1, rows, d, 0, False, False, self.te_w2, self.te_b2, emb)
        table = torch.empty(rows, self.temb_all_w.shape[0], device=self._device, dtype=torch.float32)
        ops.small_linear(emb, rows, d, 0, True, False, self.temb_all_w, self.temb_all_b, table)
        return table

    def _transformer(self, T: _TransformerParams, x: torch.Tensor, kv: torch.Tensor, ctx_len: int) -> torch.Tensor:
        K = self.K
        BF = K.op_dtype
        n, h, w, c = x.shape
        s = h * w
        m = n * s
        a = K.groupnorm(x, None, T.gn_g, T.gn_b, 1e-6, False)
        t = K.linear(a.view(m, c), T.pin_w, T.pin_b)
        ln = K.layernorm(t, *T.ln[0])
        qkv = K.linear(ln, T.qkv_w, None, out_dtype=BF)
        o = K.attention(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], n, T.heads, s, s, 0.125)
        t = K.linear(o, T.o1_w, T.o1_b, residual=t, out=t)
        ln = K.layernorm(t, *T.ln[1])
        q = K.linear(ln, T.q2_w, None, out_dtype=BF)
        o = K.attention(q, kv[:, :c], kv[:, c:], n, T.heads, s, ctx_len, 0.125)
        t = K.linear(o, T.o2_w, T.o2_b, residual=t, out=t)
        ln = K.layernorm(t, *T.ln[2])
        g = K.linear(ln, T.ff1_w, T.ff1_b, out_dtype=BF, geglu=True, block_n=T.ff_block_n)
        # (in place on the token stream when it is already in the operand format: the epilogue then adds its chunk to it
        # by a TMA reduce instead of reading it)
        t2 = K.linear(g, T.ff2_w, T.ff2_b, residual=t, out=t) if t.dtype == BF else \
            K.linear(g, T.ff2_w, T.ff2_b, residual=t, out_dtype=BF)
        # in place on the block's input (nothing else reads it): an fp32 residual that is also the output is added by a TMA
        # reduce in L2 instead of being read by the epilogue
        # (the block's output is normalised next -- by the following resnet's norm1, possibly after a detour through the
        # skip stack -- so the epilogue leaves its column sums where the tiling allows; that epilogue reads the residual)
        xr = x.view(m, c)
        o = K.linear(t2, T.pout_w, T.pout_b, residual=xr, out=xr, gn_rows_per_image=s if self._pout_sums else 0)
        return keep_tile_sums(o.view(n, h, w, c), o)

    # ---- the whole forward on NHWC tensors ---------------------------------------------------------------
    def forward_nhwc(self, latent: torch.Tensor, n_img: int, t: torch.Tensor, ctx_kv: List[torch.Tensor],
                     ctx_len: int) -> torch.Tensor:
        """latent fp32 NHWC [n_src, h, w, 4] (n_img % n_src == 0: CFG halves share a latent), t device fp32 [1] or
        [n_img]; returns fp32 NHWC [n_img, h, w, 4]."""
        K, cfg = self.K, self.config
        eps = cfg.norm_eps
        table = self.time_table(t)
        cols = self.K.im2col(latent, n_img, 1, 64)
        hw = latent.shape[1] * latent.shape[2]
        h = K.linear(cols, self.conv_in_w, self.conv_in_b, gn_rows_per_image=hw)
        skips = [h]
        kv_i = 0
        for blk in self.down:
            for j, R in enumerate(blk["res"]):
                h = run_resnet(K, R, h, None, eps, table)
                if blk["attn"] is not None:
                    h = self._transformer(blk["attn"][j], h, ctx_kv[kv_i], ctx_len)
                    kv_i += 1
                skips.append(h)
            if blk["down"] is not None:
                w, b = blk["down"]
                h = K.conv3x3(h, w, b, stride=2, gn_next=True)   # Downsample2D: 3x3, stride 2, padding 1
                skips.append(h)
        h = run_resnet(K, self.mid_res[0], h, None, eps, table)
        h = self._transformer(self.mid_attn, h, ctx_kv[kv_i], ctx_len)
        kv_i += 1
        h = run_resnet(K, self.mid_res[1], h, None, eps, table)
        for blk in self.up:
            for j, R in enumerate(blk["res"]):
                h = run_resnet(K, R, h, skips.pop(), eps, table)
                if blk["attn"] is not None:
                    h = self._transformer(blk["attn"][j], h, ctx_kv[kv_i], ctx_len)
                    kv_i += 1
            if blk["up"] is not None:
                w, b = blk["up"]
                h = K.conv3x3(K.upsample2x(h), w, b, gn_next=True)
        a = K.groupnorm(h, None, self.out_g, self.out_b, eps, True)
        return K.conv3x3(a, self.conv_out_w, self.conv_out_b, out_dtype=torch.float32)

    # ---- diffusers call signature ----------------------------------------------------------------------------
    def forward(self, sample: torch.Tensor, timestep, encoder_hidden_states: torch.Tensor, timestep_cond=None,
                cross_attention_kwargs=None, added_cond_kwargs=None, return_dict: bool = True, **_unused):
        """sample NCHW [B2,4,h,w] (any float dtype) -> NCHW noise / v prediction in the caller's dtype."""
        dev = self._device
        n = sample.shape[0]
        lat = sample.to(device=dev, dtype=torch.float32).permute(0, 2, 3, 1).contiguous()
        t = torch.as_tensor(timestep, device=dev).reshape(-1).to(torch.float32)
        if t.numel() not in (1, n):
            raise ValueError(f"timestep has {t.numel()} entries for batch {n}")
        ctx = encoder_hidden_states
        # The step-invariant cross-attention K/V are kept while the SAME tensor object is passed again, unmodified (the
        # pipeline reuses one `prompt_embeds` for all steps of a request).  The cache holds a strong reference to that
        # tensor, so its address cannot be handed to a new request's embeddings while the entry is alive.
        c = self._ctx_cache
        if c is None or c[0] is not ctx or c[1] != ctx._version:
            self._ctx_cache = c = (ctx, ctx._version, self.project_context(ctx))
        out = self.forward_nhwc(lat, n, t.contiguous(), c[2], ctx.shape[1])
        out = out.permute(0, 3, 1, 2).to(sample.dtype)
        if return_dict:
            return _Out(out)
        return (out,)


class _Out:
    def __init__(self, sample):
        self.sample = sample

    def __getitem__(self, i):
        return (self.sample,)[i]

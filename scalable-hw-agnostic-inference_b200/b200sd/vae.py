"""B200VAEDecoder -- drop-in for `pipe.vae.decoder` (+ `pipe.vae.post_quant_conv`), which the reference swaps at
app/run-sd.py:124-134 and reaches through `vae.decode(latents / scaling_factor)` inside pipe(...) (:140).

The 1x1 post_quant_conv is folded exactly into conv_in: the latent gets a fifth, all-ones channel so that the
folded bias is present at in-image taps and absent in the zero padding, as in the unfused pair.  `B200PostQuantConv`
is the identity that takes post_quant_conv's place on the pipeline object.
Mid-block attention (1 head x 512) runs on the d = 512 flash kernel (`attention512.cu`); other widths (toy
configurations) as GEMM (fp32 scores) -> row softmax -> GEMM on the tcgen05 GEMM kernel.
"""
from __future__ import annotations

from typing import Dict

import torch

from . import ops
from .config import VAEConfig
from .layers import Kernels, ResnetParams, keep_tile_sums, pack_conv3x3, run_resnet


class B200PostQuantConv(torch.nn.Module):
    """post_quant_conv lives inside B200VAEDecoder.conv_in; this keeps `vae.post_quant_conv(z)` callable."""

    def forward(self, z):
        return z


class _MidAttention:
    """the single-head, 512-wide attention of the VAE mid block (decoder and encoder share the module)"""

    def __init__(self, sd, prefix: str, dev, BF):
        f = lambda k: sd[prefix + k].to(dev).float().contiguous()
        b = lambda k: sd[prefix + k].to(dev).to(BF).contiguous()
        self.at_g, self.at_b = f("group_norm.weight"), f("group_norm.bias")
        self.wq, self.bq = b("to_q.weight"), f("to_q.bias")
        self.wk, self.bk = b("to_k.weight"), f("to_k.bias")
        self.wv, self.bv = b("to_v.weight"), f("to_v.bias")
        self.wo, self.bo = b("to_out.0.weight"), f("to_out.0.bias")
        # fused Q|K|V projection for the flash path (head dimension 512)
        self.wqkv = torch.cat([self.wq, self.wk, self.wv]).contiguous()
        self.bqkv = torch.cat([self.bq, self.bk, self.bv]).contiguous()
        import os
        self.flash = os.environ.get("B200SD_VAE_FLASH", "1") != "0"

    def __call__(self, K: Kernels, h: torch.Tensor) -> torch.Tensor:
        BF = K.op_dtype
        dev = K.device
        n, hh, ww, c = h.shape
        s = hh * ww
        a = K.groupnorm(h, None, self.at_g, self.at_b, 1e-6, False).view(n, s, c)
        if c == 512 and self.flash:
            # one head of dimension 512: the flash kernel (the score matrix never leaves the SM); all images in one launch
            qkv = K.linear(a.view(n * s, c), self.wqkv, self.bqkv, out_dtype=BF)
            o = torch.empty(n * s, c, device=dev, dtype=BF)
            ops.attention512(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], o, n, s, s, float(c) ** -0.5)
            hr = h.view(n * s, c)
            r = K.linear(o, self.wo, self.bo, residual=hr, out=hr, gn_rows_per_image=s)
            return keep_tile_sums(r.view(n, hh, ww, c), r)
        # other widths (toy configurations): GEMM (fp32 scores) -> row softmax -> GEMM, image by image
        q = K.linear(a, self.wq, self.bq, out_dtype=BF)
        k = K.linear(a, self.wk, self.bk, out_dtype=BF)
        o = torch.empty(n, s, c, device=dev, dtype=BF)
        scores = torch.empty(s, s, device=dev, dtype=torch.float32)
        probs = torch.empty(s, s, device=dev, dtype=BF)
        vt = torch.empty(c, s, device=dev, dtype=BF)
        for i in range(n):
            # V^T [c, s] = Wv [c, c] . a_i^T  (+ bv along rows): the PV GEMM then takes it as a K-major B operand
            K.linear(self.wv, a[i], self.bv, out=vt, bias_along_m=True)
            K.linear(q[i], k[i], None, out=scores)
            ops.softmax_rows(scores, probs, s, s, float(c) ** -0.5)
            K.linear(probs, vt, None, out=o[i])
        hr = h.view(n * s, c)  # in place on the block's input (dead after this)
        r = K.linear(o.view(n * s, c), self.wo, self.bo, residual=hr, out=hr, gn_rows_per_image=s)
        return keep_tile_sums(r.view(n, hh, ww, c), r)


class B200VAEDecoder(torch.nn.Module):
    def __init__(self, state_dict: Dict[str, torch.Tensor], config=None, device="cuda",
                 trunk_dtype=None, op_dtype=None, max_images_per_pass: int = 4):
        super().__init__()
        self.config = VAEConfig.from_any(config if config is not None else VAEConfig())
        cfg = self.config
        self._device = torch.device(device)
        if self._device.type != "cuda":
            raise RuntimeError("B200VAEDecoder needs a CUDA device (there is no CPU path)")
        ops.check_device(self._device.index or 0)
        self.K = Kernels(self._device, trunk_dtype, op_dtype)
        BF = self.K.op_dtype
        self.max_images_per_pass = max_images_per_pass
        dev, sd = self._device, state_dict
        f = lambda k: sd[k].to(dev).float().contiguous()
        b = lambda k: sd[k].to(dev).to(BF).contiguous()
        rev = list(reversed(cfg.block_out_channels))
        lc = cfg.latent_channels

        # conv_in o post_quant_conv, K order (tap, c') with c' = 0..lc-1 latent channels, c' = lc the ones channel
        w_in = sd["decoder.conv_in.weight"].to(dev).double()                 # [C0, lc, 3, 3]
        w_pq = sd["post_quant_conv.weight"].to(dev).double().reshape(lc, lc)  # [c, c']
        b_pq = sd["post_quant_conv.bias"].to(dev).double()
        w_taps = w_in.permute(0, 2, 3, 1)                                      # [C0, 3, 3, c]
        folded = torch.cat([w_taps @ w_pq, (w_taps @ b_pq).unsqueeze(-1)], dim=-1)  # [C0,3,3,lc+1]
        folded = folded.reshape(rev[0], 9 * (lc + 1))
        self.conv_in_w = torch.cat([folded, folded.new_zeros(rev[0], 64 - folded.shape[1])], 1).to(BF).contiguous()
        self.conv_in_b = f("decoder.conv_in.bias")

        self.mid_res = [ResnetParams(sd, f"decoder.mid_block.resnets.{j}.", dev, BF) for j in range(2)]
        self.mid_attn = _MidAttention(sd, "decoder.mid_block.attentions.0.", dev, BF)
        self.up = []
        for i, co in enumerate(rev):
            blk = {"res": [ResnetParams(sd, f"decoder.up_blocks.{i}.resnets.{j}.", dev, BF)
                           for j in range(cfg.layers_per_block + 1)], "up": None}
            if i < len(rev) - 1:
                p = f"decoder.up_blocks.{i}.upsamplers.0.conv."
                blk["up"] = (pack_conv3x3(sd[p + "weight"].to(dev), BF), f(p + "bias"))
            self.up.append(blk)
        self.out_g, self.out_b = f("decoder.conv_norm_out.weight"), f("decoder.conv_norm_out.bias")
        self.conv_out_w = pack_conv3x3(sd["decoder.conv_out.weight"].to(dev), BF)
        self.conv_out_b = f("decoder.conv_out.bias")

    @property
    def dtype(self):
        return self.K.op_dtype

    @property
    def device(self):
        return self._device

    def decode_nhwc(self, z: torch.Tensor) -> torch.Tensor:
        """z fp32 NHWC [B, h, w, 4] (already divided by scaling_factor) -> fp32 NHWC [B, 8h, 8w, 3]."""
        K = self.K
        n = z.shape[0]
        z5 = torch.cat([z, torch.ones_like(z[..., :1])], dim=-1).contiguous()
        h = K.linear(K.im2col(z5, n, 1, 64), self.conv_in_w, self.conv_in_b, gn_rows_per_image=z.shape[1] * z.shape[2])
        h = run_resnet(K, self.mid_res[0], h, None, 1e-6, None)
        h = self.mid_attn(K, h)
        h = run_resnet(K, self.mid_res[1], h, None, 1e-6, None)
        for blk in self.up:
            for R in blk["res"]:
                h = run_resnet(K, R, h, None, 1e-6, None)
            if blk["up"] is not None:
                w, b = blk["up"]
                h = K.conv3x3(K.upsample2x(h), w, b, gn_next=True)
        a = K.groupnorm(h, None, self.out_g, self.out_b, 1e-6, True)
        return K.conv3x3(a, self.conv_out_w, self.conv_out_b, out_dtype=torch.float32)

    def decode_batched_nhwc(self, z: torch.Tensor) -> torch.Tensor:
        outs = [self.decode_nhwc(z[i:i + self.max_images_per_pass].contiguous())
                for i in range(0, z.shape[0], self.max_images_per_pass)]
        return outs[0] if len(outs) == 1 else torch.cat(outs)

    def forward(self, z: torch.Tensor, *_a, **_k) -> torch.Tensor:
        """`vae.decoder(z)` signature: NCHW latent (post_quant_conv folded in) -> NCHW image in the caller's dtype."""
        zn = z.to(device=self._device, dtype=torch.float32).permute(0, 2, 3, 1).contiguous()
        img = self.decode_batched_nhwc(zn)
        return img.permute(0, 3, 1, 2).to(z.dtype)


class B200QuantConv(torch.nn.Module):
    """quant_conv lives inside B200VAEEncoder.conv_out; this keeps `vae.quant_conv(h)` callable."""

    def forward(self, h):
        return h


class B200VAEEncoder(torch.nn.Module):
    """Drop-in for `pipe.vae.encoder` (+ `pipe.vae.quant_conv`): AutoencoderKL.encode of the img2img pipelines
    (the reference loads the whole AutoencoderKL at app/run-sd.py:107).  Same kernels as the decoder:
    GroupNorm+SiLU, implicit-GEMM conv3x3, the mid-block attention; the three Downsample2D(padding=0) convolutions --
    pad (0,1,0,1), then 3x3 stride 2 -- are the conv kernel's stride-2 / pad_mode 1 form.  The 1x1 quant_conv is folded exactly into
    conv_out (it follows it with nothing in between).  Returns the moments cat(mean, logvar)."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], config=None, device="cuda", trunk_dtype=None, op_dtype=None,
                 max_images_per_pass: int = 2):
        super().__init__()
        self.config = VAEConfig.from_any(config if config is not None else VAEConfig())
        cfg = self.config
        self._device = torch.device(device)
        if self._device.type != "cuda":
            raise RuntimeError("B200VAEEncoder needs a CUDA device (there is no CPU path)")
        ops.check_device(self._device.index or 0)
        self.K = Kernels(self._device, trunk_dtype, op_dtype)
        BF = self.K.op_dtype
        self.max_images_per_pass = max_images_per_pass
        dev, sd = self._device, state_dict
        f = lambda k: sd[k].to(dev).float().contiguous()
        boc = list(cfg.block_out_channels)
        lc2 = 2 * cfg.latent_channels
        # conv_in: 3 image channels -> im2col'd GEMM with K = 27 padded to 64
        from .layers import pack_conv3x3_im2col
        self.conv_in_w = pack_conv3x3_im2col(sd["encoder.conv_in.weight"].to(dev), 64, BF)
        self.conv_in_b = f("encoder.conv_in.bias")
        self.down = []
        for i, _co in enumerate(boc):
            blk = {"res": [ResnetParams(sd, f"encoder.down_blocks.{i}.resnets.{j}.", dev, BF)
                           for j in range(cfg.layers_per_block)], "down": None}
            if i < len(boc) - 1:
                p = f"encoder.down_blocks.{i}.downsamplers.0.conv."
                blk["down"] = (pack_conv3x3(sd[p + "weight"].to(dev), BF), f(p + "bias"))
            self.down.append(blk)
        self.mid_res = [ResnetParams(sd, f"encoder.mid_block.resnets.{j}.", dev, BF) for j in range(2)]
        self.mid_attn = _MidAttention(sd, "encoder.mid_block.attentions.0.", dev, BF)
        self.out_g, self.out_b = f("encoder.conv_norm_out.weight"), f("encoder.conv_norm_out.bias")
        # quant_conv o conv_out: W'[o, c, ky, kx] = sum_m Wq[o, m] W[m, c, ky, kx];  b' = Wq b + bq
        w_out = sd["encoder.conv_out.weight"].to(dev).double()                  # [2lc, C, 3, 3]
        w_q = sd["quant_conv.weight"].to(dev).double().reshape(lc2, lc2)
        folded = torch.einsum("om,mcyx->ocyx", w_q, w_out)
        self.conv_out_w = pack_conv3x3(folded.float(), BF)
        self.conv_out_b = (w_q @ sd["encoder.conv_out.bias"].to(dev).double()
                           + sd["quant_conv.bias"].to(dev).double()).float().contiguous()

    @property
    def dtype(self):
        return self.K.op_dtype

    @property
    def device(self):
        return self._device

    def encode_nhwc(self, x: torch.Tensor) -> torch.Tensor:
        """x fp32 NHWC [B, H, W, 3] in [-1, 1] -> moments fp32 NHWC [B, H/8, W/8, 2*latent]"""
        K = self.K
        n, hh, ww, _ = x.shape
        h = K.linear(K.im2col(x.contiguous(), n, 1, 64), self.conv_in_w, self.conv_in_b, gn_rows_per_image=hh * ww)
        for blk in self.down:
            for R in blk["res"]:
                h = run_resnet(K, R, h, None, 1e-6, None)
            if blk["down"] is not None:
                w, b = blk["down"]
                h = K.conv3x3(h, w, b, stride=2, pad_mode=1, gn_next=True)  # pad (0,1,0,1), 3x3 stride 2, no padding
        h = run_resnet(K, self.mid_res[0], h, None, 1e-6, None)
        h = self.mid_attn(K, h)
        h = run_resnet(K, self.mid_res[1], h, None, 1e-6, None)
        a = K.groupnorm(h, None, self.out_g, self.out_b, 1e-6, True)
        return K.conv3x3(a, self.conv_out_w, self.conv_out_b, out_dtype=torch.float32)

    def forward(self, x: torch.Tensor, *_a, **_k) -> torch.Tensor:
        """`vae.encoder(x)` signature: NCHW image in [-1, 1] -> NCHW moments (quant_conv folded in), caller's dtype"""
        xn = x.to(device=self._device, dtype=torch.float32).permute(0, 2, 3, 1).contiguous()
        outs = [self.encode_nhwc(xn[i:i + self.max_images_per_pass])
                for i in range(0, xn.shape[0], self.max_images_per_pass)]
        m = outs[0] if len(outs) == 1 else torch.cat(outs)
        return m.permute(0, 3, 1, 2).to(x.dtype)

    def encode_latents(self, x: torch.Tensor, noise: torch.Tensor = None) -> torch.Tensor:
        """what img2img feeds the denoising loop: scaling_factor * (mean + exp(0.5 * clamp(logvar, -30, 20)) * noise),
        or scaling_factor * mean (the distribution's mode) without noise"""
        m = self.forward(x).float()
        mean, logvar = torch.chunk(m, 2, dim=1)
        z = mean if noise is None else mean + torch.exp(0.5 * logvar.clamp(-30.0, 20.0)) * noise.to(m.device)
        return self.config.scaling_factor * z

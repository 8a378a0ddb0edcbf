"""B200VAEDecoder -- drop-in for `pipe.vae.decoder` (+ `pipe.vae.post_quant_conv`), which the reference swaps at
app/run-sd.py:124-134 and reaches through `vae.decode(latents / scaling_factor)` inside pipe(...) (:140).

The 1x1 post_quant_conv is folded exactly into conv_in: the latent gets a fifth, all-ones channel so that the
folded bias is present at in-image taps and absent in the zero padding, as in the unfused pair.  `B200PostQuantConv`
is the identity that takes post_quant_conv's place on the pipeline object.
Mid-block attention (1 head x 512) runs as GEMM (fp32 scores) -> row softmax -> GEMM on the same tcgen05 kernel.
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
        a = "decoder.mid_block.attentions.0."
        self.at_g, self.at_b = f(a + "group_norm.weight"), f(a + "group_norm.bias")
        self.wq, self.bq = b(a + "to_q.weight"), f(a + "to_q.bias")
        self.wk, self.bk = b(a + "to_k.weight"), f(a + "to_k.bias")
        self.wv, self.bv = b(a + "to_v.weight"), f(a + "to_v.bias")
        self.wo, self.bo = b(a + "to_out.0.weight"), f(a + "to_out.0.bias")
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

    def _mid_attention(self, h: torch.Tensor) -> torch.Tensor:
        K = self.K
        BF = K.op_dtype
        n, hh, ww, c = h.shape
        s = hh * ww
        a = K.groupnorm(h, None, self.at_g, self.at_b, 1e-6, False).view(n, s, c)
        q = K.linear(a, self.wq, self.bq, out_dtype=BF)
        k = K.linear(a, self.wk, self.bk, out_dtype=BF)
        o = torch.empty(n, s, c, device=self._device, dtype=BF)
        scores = torch.empty(s, s, device=self._device, dtype=torch.float32)
        probs = torch.empty(s, s, device=self._device, dtype=BF)
        vt = torch.empty(c, s, device=self._device, dtype=BF)
        for i in range(n):
            # V^T [c, s] = Wv [c, c] . a_i^T  (+ bv along rows): the PV GEMM then takes it as a K-major B operand
            K.linear(self.wv, a[i], self.bv, out=vt, bias_along_m=True)
            K.linear(q[i], k[i], None, out=scores)
            ops.softmax_rows(scores, probs, s, s, float(c) ** -0.5)
            K.linear(probs, vt, None, out=o[i])
        hr = h.view(n * s, c)  # in place on the block's input (dead after this)
        r = K.linear(o.view(n * s, c), self.wo, self.bo, residual=hr, out=hr, gn_rows_per_image=s)
        return keep_tile_sums(r.view(n, hh, ww, c), r)

    def decode_nhwc(self, z: torch.Tensor) -> torch.Tensor:
        """z fp32 NHWC [B, h, w, 4] (already divided by scaling_factor) -> fp32 NHWC [B, 8h, 8w, 3]."""
        K = self.K
        n = z.shape[0]
        z5 = torch.cat([z, torch.ones_like(z[..., :1])], dim=-1).contiguous()
        h = K.linear(K.im2col(z5, n, 1, 64), self.conv_in_w, self.conv_in_b, gn_rows_per_image=z.shape[1] * z.shape[2])
        h = run_resnet(K, self.mid_res[0], h, None, 1e-6, None)
        h = self._mid_attention(h)
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

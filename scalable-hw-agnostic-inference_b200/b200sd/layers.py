"""Weight packing and the per-layer call helpers shared by B200UNet and B200VAEDecoder.

Layouts: activations NHWC; every tensor-core operand 16-bit in `op_dtype` (fp16 by default,
bf16 selectable -- see DESIGN.md "operand format"); the residual trunk in `trunk_dtype` (fp16 by default, fp32 / bf16 selectable);
conv weights KRSC [Cout, 9*Cin]; linear weights [out, in].
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import torch

from . import ops

_OP_DTYPES = {"fp16": torch.float16, "float16": torch.float16, "bf16": torch.bfloat16, "bfloat16": torch.bfloat16}


def default_op_dtype() -> torch.dtype:
    """16-bit format of tensor-core operands.  fp16 keeps the single-step UNet error vs the fp32 oracle at ~1e-3;
    bf16 (the reference's storage dtype, app/run-sd.py:39) sits at ~1.1e-2 with identical tensor throughput."""
    return _OP_DTYPES[os.environ.get("B200SD_OPERAND_DTYPE", "fp16").lower()]


def default_trunk_dtype() -> torch.dtype:
    """dtype of the residual stream (the tensors every block adds into).  fp16 by default: the reference keeps it in
    bf16 (its only dtype, app/run-sd.py:39); fp16 has three more significand bits, every sum is still formed in fp32
    inside the epilogue and rounded once, and stores saturate instead of overflowing.  Measured against the fp32 oracle at
    768^2: single-step rel-L2 1.7e-3 with the fp16 trunk, 1.3e-3 with an fp32 trunk (B200SD_TRUNK_DTYPE=fp32), both far
    inside the 1e-2 bound, for 3.5 % more images/s (the trunk is read and written by every memory-bound kernel)."""
    name = os.environ.get("B200SD_TRUNK_DTYPE", "fp16").lower()
    return torch.float32 if name in ("fp32", "float32") else _OP_DTYPES[name]


# ------------------------------------------------------------------------------------------------ packing
def pack_conv3x3(w: torch.Tensor, dtype: torch.dtype = torch.float16) -> torch.Tensor:
    """[Cout, Cin, 3, 3] -> 16-bit [Cout, 9*Cin], K order (ky, kx, c)."""
    co, ci = w.shape[:2]
    return w.permute(0, 2, 3, 1).reshape(co, 9 * ci).to(dtype).contiguous()


def pack_conv3x3_im2col(w: torch.Tensor, k_pad: int, dtype: torch.dtype = torch.float16) -> torch.Tensor:
    p = pack_conv3x3(w, dtype)
    if p.shape[1] < k_pad:
        p = torch.cat([p, p.new_zeros(p.shape[0], k_pad - p.shape[1])], dim=1)
    return p.contiguous()


def pack_conv1x1(w: torch.Tensor, dtype: torch.dtype = torch.float16) -> torch.Tensor:
    return w.reshape(w.shape[0], w.shape[1]).to(dtype).contiguous()


def pack_geglu(w: torch.Tensor, b: torch.Tensor, block_n: int,
               dtype: torch.dtype = torch.float16) -> Tuple[torch.Tensor, torch.Tensor]:
    """[2*inner, K] (values then gates) -> rows interleaved per block_n tile: tile t = values[t*h:(t+1)*h] ++ gates[...]"""
    inner = w.shape[0] // 2
    h = block_n // 2
    assert inner % h == 0
    idx = torch.arange(inner, device=w.device).view(-1, h)
    order = torch.cat([idx, idx + inner], dim=1).reshape(-1)
    return w[order].to(dtype).contiguous(), b[order].float().contiguous()


# ------------------------------------------------------------------------------------------------ call helpers
def tile_sums_of(x: torch.Tensor) -> Optional[torch.Tensor]:
    """the column-sum table the kernel that wrote `x` attached to it (Kernels.conv3x3 / linear with gn_next), if any.
    It rides on the tensor OBJECT, so a tensor that merely reuses the address later never inherits it; code that takes a
    view of such a tensor and passes the view on copies it with `keep_tile_sums`."""
    return getattr(x, "_b200_tile_sums", None)


def keep_tile_sums(view: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    t = tile_sums_of(src)
    if t is not None:
        view._b200_tile_sums = t
    return view


class Kernels:
    """Helper bound to one device: owns the GroupNorm scratch and the dtype policy."""

    def __init__(self, device: torch.device, trunk_dtype: Optional[torch.dtype] = None,
                 op_dtype: Optional[torch.dtype] = None):
        self.device = device
        self.trunk_dtype = trunk_dtype or default_trunk_dtype()
        self.op_dtype = op_dtype or default_op_dtype()
        assert self.op_dtype in (torch.float16, torch.bfloat16)
        self._gn_ws: Optional[torch.Tensor] = None
        # GroupNorm statistics from the producing kernel's epilogue instead of a pass over the tensor (B200SD_GN_FUSE=0: off)
        self.fuse_gn_stats = os.environ.get("B200SD_GN_FUSE", "1") != "0"
        # split-K scratch for GEMMs with fewer output tiles than SMs; zero-filled once, the kernel keeps its counters clean
        self.gemm_ws = torch.zeros(int(os.environ.get("B200SD_GEMM_WS_MB", "96")) << 18, device=device,
                                   dtype=torch.float32) if device.type == "cuda" else None

    def gn_workspace(self, n_img: int, hw: int, groups: int) -> torch.Tensor:
        need = ops.groupnorm_workspace_bytes(n_img, hw, groups)
        if self._gn_ws is None or self._gn_ws.numel() * 4 < need:
            self._gn_ws = torch.empty(max(need // 4, 1 << 18), device=self.device, dtype=torch.float32)
        return self._gn_ws

    def groupnorm(self, x1: torch.Tensor, x2: Optional[torch.Tensor], gamma, beta, eps: float, silu: bool,
                  want_raw: bool = False, groups: int = 32):
        """GroupNorm(+SiLU) of x1 (or of the channel concat [x1, x2]).  When the kernels that produced x1 (and x2) left
        their column sums behind (`tile_sums_of`), the statistics pass over the tensors is skipped: a small kernel reduces
        those tables to per-(image, group) sums and only the normalising pass reads the activations."""
        n, h, w, c1 = x1.shape
        c = c1 + (x2.shape[-1] if x2 is not None else 0)
        y = torch.empty(n, h, w, c, device=self.device, dtype=self.op_dtype)
        raw = torch.empty_like(y) if want_raw else None
        t1 = tile_sums_of(x1)
        t2 = tile_sums_of(x2) if x2 is not None else None
        if self.fuse_gn_stats and t1 is not None and (x2 is None or t2 is not None):
            sums = torch.empty(n, groups, 2, device=self.device, dtype=torch.float32)
            ops.gn_tile_sums(t1, t2, n, groups, sums)
            ops.groupnorm(x1, x2, n, h * w, groups, eps, gamma, beta, silu, y, raw, None, sums, 0, None)
        else:
            ops.groupnorm(x1, x2, n, h * w, groups, eps, gamma, beta, silu, y, raw, self.gn_workspace(n, h * w, groups))
        return (y, raw) if want_raw else y

    def _tile_sums_buffer(self, gemm_args, n_img: int, cout: int, rows_per_image: int = 0) -> Optional[torch.Tensor]:
        if not self.fuse_gn_stats:
            return None
        parts = ops.gemm_tile_sum_parts(*gemm_args, rows_per_image=rows_per_image)
        if parts <= 0:
            return None
        return torch.empty(n_img, parts, cout, 2, device=self.device, dtype=torch.float32)

    def layernorm(self, x: torch.Tensor, gamma, beta, eps: float = 1e-5) -> torch.Tensor:
        y = torch.empty(x.shape, device=self.device, dtype=self.op_dtype)
        ops.layernorm(x, eps, gamma, beta, y)
        return y

    def conv3x3(self, a: torch.Tensor, w: torch.Tensor, bias, *, row_bias=None, residual=None,
                out_dtype=None, gn_next: bool = False, stride: int = 1, pad_mode: int = 0) -> torch.Tensor:
        """a 16-bit NHWC, w packed [Cout, 9*Cin].  stride 1 or 2 (Downsample2D: the TMA box takes every second pixel, no
        im2col copy); pad_mode 0 = Conv2d(padding=1), 1 = pad (0,1,0,1) and no padding (the VAE encoder's Downsample2D).
        gn_next: the output feeds a GroupNorm -- have the epilogue emit its column sums (when this problem's tiling
        allows it) and attach them to the result."""
        if a.dtype != self.op_dtype:
            # a trunk tensor fed straight to a convolution (Downsample2D) while the trunk is kept in another format than the
            # tensor-core operands (B200SD_TRUNK_DTYPE=fp32 / bf16 with fp16 operands): one conversion pass
            a16 = torch.empty(a.shape, device=self.device, dtype=self.op_dtype)
            ops.cast(a.contiguous(), a16)
            a = a16
        n, h, wd, cin = a.shape
        cout = w.shape[0]
        pad_lo = 0 if pad_mode == 1 else 1
        oh, ow = ((h + pad_lo - 2) // stride + 1, (wd + pad_lo - 2) // stride + 1) if (stride == 2 or pad_mode) else (h, wd)
        out = torch.empty(n, oh, ow, cout, device=self.device, dtype=out_dtype or self.trunk_dtype)
        rb_stride = 0
        if row_bias is not None and row_bias.shape[0] > 1:
            rb_stride = row_bias.stride(0)
        res2 = residual.view(-1, cout) if residual is not None else None
        args = (a, n, h, wd, cin, 9, w, bias, False, row_bias, rb_stride, res2, out.view(-1, cout), False, 0)
        conv = {"conv_stride": stride, "conv_pad_mode": pad_mode}
        ts = None
        if gn_next and self.fuse_gn_stats:
            parts = ops.gemm_tile_sum_parts(*args, **conv)
            if parts > 0:
                ts = torch.empty(n, parts, cout, 2, device=self.device, dtype=torch.float32)
        if ts is not None:
            ops.gemm(*args, None, 1, 0, ts, 0, stride, pad_mode)
            out._b200_tile_sums = ts
        else:
            ops.gemm(*args, self.gemm_ws, 0, 0, None, 0, stride, pad_mode)
        return out

    def linear(self, a: torch.Tensor, w: torch.Tensor, bias, *, residual=None, out_dtype=None, geglu=False,
               out: Optional[torch.Tensor] = None, bias_along_m: bool = False, block_n: int = 0,
               split_k: int = 0, gelu: bool = False, gn_rows_per_image: int = 0) -> torch.Tensor:
        """a 16-bit [M, K] (any leading dims), w 16-bit [N, K].  gelu: erf-GELU on (a w^T + bias), before the residual.
        gn_rows_per_image > 0: the output is an NHWC activation of that many pixels per image and feeds a GroupNorm --
        emit the column sums (see conv3x3)."""
        k = a.shape[-1]
        m = a.numel() // k
        n_cols = w.shape[0] // 2 if geglu else w.shape[0]
        if out is None:
            out = torch.empty(*a.shape[:-1], n_cols, device=self.device, dtype=out_dtype or self.trunk_dtype)
        res2 = residual.view(m, n_cols) if residual is not None else None
        args = (a, 1, 1, m, k, 1, w, bias, bias_along_m, None, 0, res2, out.view(m, n_cols), geglu, block_n)
        ts = None
        if gn_rows_per_image > 0 and split_k <= 1 and m % gn_rows_per_image == 0:
            ts = self._tile_sums_buffer(args, m // gn_rows_per_image, n_cols, gn_rows_per_image)
        if ts is not None:
            ops.gemm(*args, None, 1, 1 if gelu else 0, ts, gn_rows_per_image)
            out._b200_tile_sums = ts
        else:
            ops.gemm(*args, self.gemm_ws, split_k, 1 if gelu else 0)
        return out

    def im2col(self, x: torch.Tensor, n_img: int, stride: int, k_pad: int, pad_lo: int = 1) -> torch.Tensor:
        """pad_lo = 1: Conv2d(padding=1); pad_lo = 0: pad (0,1,0,1) then padding 0 (the VAE encoder's Downsample2D)"""
        n_src, h, w, cin = x.shape
        oh, ow = (h + pad_lo - 2) // stride + 1, (w + pad_lo - 2) // stride + 1
        out = torch.empty(n_img, oh, ow, k_pad, device=self.device, dtype=self.op_dtype)
        ops.im2col3x3(x, n_img, n_src, h, w, cin, stride, out, pad_lo)
        return out

    def upsample2x(self, x: torch.Tensor) -> torch.Tensor:
        n, h, w, c = x.shape
        out = torch.empty(n, 2 * h, 2 * w, c, device=self.device, dtype=self.op_dtype)
        ops.upsample2x(x, n, h, w, c, out)
        return out

    def attention(self, q, k, v, batch: int, heads: int, sq: int, skv: int, scale: float,
                  causal: bool = False) -> torch.Tensor:
        out = torch.empty(batch * sq, heads * 64, device=self.device, dtype=q.dtype)
        ops.attention(q, k, v, out, batch, heads, sq, skv, scale, causal)
        return out


class ResnetParams:
    """packed weights of one ResnetBlock2D (keys: diffusers `...resnets.j.*`)."""

    def __init__(self, sd: Dict[str, torch.Tensor], prefix: str, dev, dtype: torch.dtype,
                 temb_slice: Optional[Tuple[int, int]] = None):
        g = lambda k: sd[prefix + k].to(dev)
        self.n1_g, self.n1_b = g("norm1.weight").float().contiguous(), g("norm1.bias").float().contiguous()
        self.n2_g, self.n2_b = g("norm2.weight").float().contiguous(), g("norm2.bias").float().contiguous()
        self.w1 = pack_conv3x3(g("conv1.weight"), dtype)
        self.b1 = g("conv1.bias").float().contiguous()
        self.w2 = pack_conv3x3(g("conv2.weight"), dtype)
        self.b2 = g("conv2.bias").float().contiguous()
        self.cout = self.w1.shape[0]
        self.has_shortcut = (prefix + "conv_shortcut.weight") in sd
        if self.has_shortcut:
            self.ws = pack_conv1x1(g("conv_shortcut.weight"), dtype)
            self.bs = g("conv_shortcut.bias").float().contiguous()
        self.temb_slice = temb_slice  # columns of the per-step timestep table (conv1 bias folded in)


def run_resnet(K: Kernels, P: ResnetParams, x: torch.Tensor, skip: Optional[torch.Tensor], eps: float,
               temb_table: Optional[torch.Tensor]) -> torch.Tensor:
    """GN+SiLU -> conv3x3 (+temb) -> GN+SiLU -> conv3x3 + shortcut; `skip` is concatenated on channels."""
    if P.has_shortcut:
        a, raw = K.groupnorm(x, skip, P.n1_g, P.n1_b, eps, True, want_raw=True)
    else:
        assert skip is None
        a, raw = K.groupnorm(x, None, P.n1_g, P.n1_b, eps, True), None
    if temb_table is not None:
        lo, hi = P.temb_slice
        h1 = K.conv3x3(a, P.w1, None, row_bias=temb_table[:, lo:hi], out_dtype=K.op_dtype, gn_next=True)
    else:
        h1 = K.conv3x3(a, P.w1, P.b1, out_dtype=K.op_dtype, gn_next=True)
    b = K.groupnorm(h1, None, P.n2_g, P.n2_b, eps, True)
    if P.has_shortcut:
        res = K.linear(raw, P.ws, P.bs)  # 1x1 conv == linear over pixels
    else:
        res = x
    # every resnet output is normalised next (a transformer's GroupNorm, the next resnet's norm1, conv_norm_out)
    return K.conv3x3(b, P.w2, P.b2, residual=res, gn_next=True)

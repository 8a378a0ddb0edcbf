"""Patch-parallel denoising of ONE large image over the GPUs of a box (BASELINE.json configs[4]: SD2.1 1024^2, a single
image split across 8 B200).  The reference has no counterpart (one GPU per pod,
sd21-a10g-cuda-deploy.yaml:47-49); north_star names the mode as optional, with NCCL over NVLink all-gathering the
attention K/V.

Partition: latent ROWS.  Rank r of W owns rows [r*H/W, (r+1)*H/W) of every feature map (H = 128 / 64 / 32 / 16 at
1024^2), all channels, the whole CFG batch; every rank holds all weights.  What crosses NVLink, per UNet step:
  * conv3x3 (stride 1 and 2): one halo row from each neighbour (NCCL send/recv), zero rows at the image's top and bottom;
  * GroupNorm: the per-(image, group) sum / sum of squares of the local rows, all-reduced (64 floats per image);
  * self-attention: the local K|V rows, all-gathered into a LOCAL buffer that the attention kernel then reads through
    tensor maps over this GPU's memory only (TMA loads here read only this GPU's memory); the gather is issued on a side
    stream right after the K|V projection and overlaps the Q projection;
  * nothing for cross-attention, LayerNorm, the feed-forward, 1x1 convolutions, up-sampling, the scheduler step.
Every kernel is the single-GPU one (libb200sd); only the orchestration differs.  The serving default (`bench.py --gpus N`,
replicas without a collective) does not touch this module.

`BandComm` holds the partition arithmetic and the three exchanges on torch.distributed tensors; it runs on gloo with
CPU tensors too, which is how tests/test_multi_cpu.py checks the index logic without a GPU.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


class BandComm:
    """row-band partition of NHWC feature maps over the ranks of a process group + the exchanges the UNet needs"""

    def __init__(self, group=None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.bytes_gathered = 0      # received by all_gather_rows (K|V), for the bench line
        self.bytes_halo = 0          # received halo rows
        self.n_allreduce = 0

    # ---- partition ------------------------------------------------------------------------------------------------
    def band(self, height: int) -> Tuple[int, int]:
        """(first row, row count) of this rank; heights must divide evenly (SD2.1 sizes do down to 16 rows over 8 ranks)"""
        if height % self.world:
            raise ValueError(f"{height} rows do not split over {self.world} ranks")
        rows = height // self.world
        return self.rank * rows, rows

    # ---- halo rows -----------------------------------------------------------------------------------------------
    def halo_pad(self, x: torch.Tensor, left_pad: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x [n, r, w, c] (this rank's band) -> [n, r + 2, w + left_pad, c]: row 0 is the upper neighbour's last row, row
        r + 1 the lower neighbour's first row (zeros at the image's top / bottom edge); `left_pad` zero columns are put in
        front of every row (the stride-2 convolution's left padding)."""
        n, r, w, c = x.shape
        if out is None:
            out = torch.empty(n, r + 2, w + left_pad, c, device=x.device, dtype=x.dtype)
        if left_pad:
            out[:, :, :left_pad].zero_()
        out[:, 1:r + 1, left_pad:].copy_(x)
        up, down = self.rank - 1, self.rank + 1
        first, last = x[:, 0].contiguous(), x[:, r - 1].contiguous()
        top, bot = torch.empty_like(first), torch.empty_like(last)
        ops = []
        if up >= 0:
            ops += [dist.P2POp(dist.isend, first, self._global(up), self.group),
                    dist.P2POp(dist.irecv, top, self._global(up), self.group)]
        if down < self.world:
            ops += [dist.P2POp(dist.isend, last, self._global(down), self.group),
                    dist.P2POp(dist.irecv, bot, self._global(down), self.group)]
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        if up >= 0:
            out[:, 0, left_pad:].copy_(top)
            self.bytes_halo += top.numel() * top.element_size()
        else:
            out[:, 0].zero_()
        if down < self.world:
            out[:, r + 1, left_pad:].copy_(bot)
            self.bytes_halo += bot.numel() * bot.element_size()
        else:
            out[:, r + 1].zero_()
        return out

    def _global(self, group_rank: int) -> int:
        return group_rank if self.group is None else dist.get_global_rank(self.group, group_rank)

    # ---- statistics ---------------------------------------------------------------------------------------------
    def allreduce_sums(self, sums: torch.Tensor) -> torch.Tensor:
        """per-(image, group) (sum, sum of squares) over the local rows -> over the whole image, in place"""
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=self.group)
        self.n_allreduce += 1
        return sums

    # ---- K|V rows ------------------------------------------------------------------------------------------------
    def all_gather_rows(self, local: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """local [n, s_local, c] (this rank's tokens of every image) -> [n, world * s_local, c] in rank order = row order,
        written into a buffer in THIS device's memory.  One gather per image, so each lands contiguously where the attention
        kernel's [batch * skv, c] view expects it."""
        n, s_local, c = local.shape
        if out is None:
            out = torch.empty(n, self.world * s_local, c, device=local.device, dtype=local.dtype)
        for b in range(n):
            dist.all_gather_into_tensor(out[b].view(-1), local[b].contiguous().view(-1), group=self.group)
        self.bytes_gathered += (self.world - 1) * local.numel() * local.element_size()
        return out

    def all_gather_bands(self, band: torch.Tensor) -> torch.Tensor:
        """[n, r, w, c] bands of all ranks -> the whole [n, world * r, w, c] tensor on every rank"""
        n, r, w, c = band.shape
        return self.all_gather_rows(band.reshape(n, r * w, c)).view(n, self.world * r, w, c)


class PatchParallelUNet:
    """B200UNet.forward_nhwc on a row band.  Shares the wrapped UNet's packed weights and kernels helper; activations are
    the band's rows only, so per-GPU activation memory and FLOPs are 1/world of the single-GPU step (plus halo rows)."""

    def __init__(self, unet, comm: BandComm):
        self.unet, self.comm = unet, comm
        self.K = unet.K
        self._side = torch.cuda.Stream(device=unet.device)   # K|V all-gather runs here, beside the Q projection

    # ---- building blocks ------------------------------------------------------------------------------------------
    def _groupnorm(self, x1, x2, gamma, beta, eps: float, silu: bool, want_raw: bool = False):
        from . import ops
        K = self.K
        n, r, w, c1 = x1.shape
        c = c1 + (x2.shape[-1] if x2 is not None else 0)
        sums = torch.empty(n, 32, 2, device=K.device, dtype=torch.float32)
        ops.groupnorm(x1, x2, n, r * w, 32, eps, gamma, beta, silu, None, None, K.gn_workspace(n, r * w, 32), None, 0, sums)
        self.comm.allreduce_sums(sums)
        y = torch.empty(n, r, w, c, device=K.device, dtype=K.op_dtype)
        raw = torch.empty_like(y) if want_raw else None
        ops.groupnorm(x1, x2, n, r * w, 32, eps, gamma, beta, silu, y, raw, None, sums, r * w * self.comm.world, None)
        return (y, raw) if want_raw else y

    def _conv3x3(self, a, w, bias, *, row_bias=None, residual=None, out_dtype=None):
        """stride-1 3x3 convolution of a band: the band is extended by its two halo rows, the single-GPU kernel runs on the
        extended band with its own zero padding, and the two outer output rows (which saw that padding instead of real
        neighbours) are dropped.  The residual rides along in the same extended layout so the epilogue still adds it."""
        K = self.K
        n, r, wd, _ = a.shape
        ap = self.comm.halo_pad(a)
        resp = None
        if residual is not None:
            resp = torch.empty(n, r + 2, wd, residual.shape[-1], device=K.device, dtype=residual.dtype)
            resp[:, 1:r + 1].copy_(residual)
            resp[:, 0].zero_()
            resp[:, r + 1].zero_()
        out = K.conv3x3(ap, w, bias, row_bias=row_bias, residual=resp, out_dtype=out_dtype)
        return out[:, 1:r + 1].contiguous()

    def _conv3x3_stride2(self, a, w, bias):
        """Downsample2D: output row R reads input rows 2R-1..2R+1, i.e. rows 2i..2i+2 of the halo-extended band, and input
        columns 2x-1..2x+1, i.e. columns 2x..2x+2 after one zero column is put in front: the conv kernel's stride-2 form
        with pad_mode 1 (nothing before, one zero row / column after).  The last output row lies past the band."""
        n, r, wd, c = a.shape
        ap = self.comm.halo_pad(a, left_pad=1)
        out = self.K.conv3x3(ap, w, bias, stride=2, pad_mode=1)      # [n, r/2 + 1, wd/2, cout]
        return out[:, : r // 2].contiguous()

    def _conv_in(self, latent_band, n_img: int):
        K, u = self.K, self.unet
        lp = self.comm.halo_pad(latent_band)
        n_src, rp, wd, _ = lp.shape
        cols = K.im2col(lp, n_img, 1, 64)                    # rows 0 and rp-1 are discarded below
        return K.linear(cols[:, 1:rp - 1].contiguous(), u.conv_in_w, u.conv_in_b)

    def _resnet(self, P, x, skip, eps: float, table):
        K = self.K
        if P.has_shortcut:
            a, raw = self._groupnorm(x, skip, P.n1_g, P.n1_b, eps, True, want_raw=True)
        else:
            a, raw = self._groupnorm(x, None, P.n1_g, P.n1_b, eps, True), None
        lo, hi = P.temb_slice
        h1 = self._conv3x3(a, P.w1, None, row_bias=table[:, lo:hi], out_dtype=K.op_dtype)
        b = self._groupnorm(h1, None, P.n2_g, P.n2_b, eps, True)
        res = K.linear(raw, P.ws, P.bs) if P.has_shortcut else x
        return self._conv3x3(b, P.w2, P.b2, residual=res)

    def _transformer(self, T, x, kv_ctx, ctx_len: int):
        K, comm = self.K, self.comm
        BF = K.op_dtype
        n, r, w, c = x.shape
        s_local = r * w
        m = n * s_local
        a = self._groupnorm(x, None, T.gn_g, T.gn_b, 1e-6, False)
        t = K.linear(a.view(m, c), T.pin_w, T.pin_b)
        ln = K.layernorm(t, *T.ln[0])
        # K|V of the local rows first, so that their gather is in flight while Q is projected
        kv_local = K.linear(ln, T.qkv_w[c:], None, out_dtype=BF)                     # [m, 2c]
        cur = torch.cuda.current_stream()
        # [n, S, 2c] in local memory; allocated on the compute stream (which joins the side stream before the buffer is
        # read or freed), filled on the side stream
        kv_all = torch.empty(n, comm.world * s_local, 2 * c, device=K.device, dtype=BF)
        self._side.wait_stream(cur)
        with torch.cuda.stream(self._side):
            comm.all_gather_rows(kv_local.view(n, s_local, 2 * c), out=kv_all)
        q = K.linear(ln, T.qkv_w[:c], None, out_dtype=BF)
        cur.wait_stream(self._side)
        kv2 = kv_all.view(n * comm.world * s_local, 2 * c)
        o = K.attention(q, kv2[:, :c], kv2[:, c:], n, T.heads, s_local, comm.world * s_local, 0.125)
        t = K.linear(o, T.o1_w, T.o1_b, residual=t, out=t)
        ln = K.layernorm(t, *T.ln[1])
        q = K.linear(ln, T.q2_w, None, out_dtype=BF)
        o = K.attention(q, kv_ctx[:, :c], kv_ctx[:, c:], n, T.heads, s_local, ctx_len, 0.125)
        t = K.linear(o, T.o2_w, T.o2_b, residual=t, out=t)
        ln = K.layernorm(t, *T.ln[2])
        g = K.linear(ln, T.ff1_w, T.ff1_b, out_dtype=BF, geglu=True, block_n=T.ff_block_n)
        # (in place on the token stream when it is already in the operand format: the epilogue then adds its chunk to it
        # by a TMA reduce instead of reading it)
        t2 = K.linear(g, T.ff2_w, T.ff2_b, residual=t, out=t) if t.dtype == BF else \
            K.linear(g, T.ff2_w, T.ff2_b, residual=t, out_dtype=BF)
        xr = x.view(m, c)
        return K.linear(t2, T.pout_w, T.pout_b, residual=xr, out=xr).view(n, r, w, c)

    # ---- the forward on a band ------------------------------------------------------------------------------------
    @torch.no_grad()
    def forward_band(self, latent_band: torch.Tensor, n_img: int, t: torch.Tensor, ctx_kv: List[torch.Tensor],
                     ctx_len: int) -> torch.Tensor:
        """latent_band fp32 NHWC [n_src, rows, w, 4] (this rank's rows; n_img % n_src == 0), t device fp32 [1] or [n_img]
        -> fp32 NHWC [n_img, rows, w, 4]: this rank's rows of the noise / v prediction."""
        u, K = self.unet, self.K
        eps = u.config.norm_eps
        table = u.time_table(t)
        h = self._conv_in(latent_band, n_img)
        skips = [h]
        kv_i = 0
        for blk in u.down:
            for j, R in enumerate(blk["res"]):
                h = self._resnet(R, h, None, eps, table)
                if blk["attn"] is not None:
                    h = self._transformer(blk["attn"][j], h, ctx_kv[kv_i], ctx_len)
                    kv_i += 1
                skips.append(h)
            if blk["down"] is not None:
                w, b = blk["down"]
                h = self._conv3x3_stride2(h, w, b)
                skips.append(h)
        h = self._resnet(u.mid_res[0], h, None, eps, table)
        h = self._transformer(u.mid_attn, h, ctx_kv[kv_i], ctx_len)
        kv_i += 1
        h = self._resnet(u.mid_res[1], h, None, eps, table)
        for blk in u.up:
            for j, R in enumerate(blk["res"]):
                h = self._resnet(R, h, skips.pop(), eps, table)
                if blk["attn"] is not None:
                    h = self._transformer(blk["attn"][j], h, ctx_kv[kv_i], ctx_len)
                    kv_i += 1
            if blk["up"] is not None:
                w, b = blk["up"]
                h = self._conv3x3(K.upsample2x(h), w, b)
        a = self._groupnorm(h, None, u.out_g, u.out_b, eps, True)
        return self._conv3x3(a, u.conv_out_w, u.conv_out_b, out_dtype=torch.float32)


class PatchParallelPipeline:
    """One image (or a small batch) denoised by all ranks together: every rank calls this with the SAME latents and
    embeddings; each denoises its band (UNet band forward + the fused CFG / scheduler kernel on the band), the final
    latent bands are all-gathered and every rank decodes the whole image (the VAE decode is 5 % of an image; it is not
    split).  The step -- ~900 kernel launches, ~140 NCCL calls on two streams, the halo copies -- is captured into ONE CUDA
    graph per rank and replayed for every step (timestep and step index live in device memory, as in B200Pipeline);
    eager, the step is bound by the host's launch rate, not by the GPUs."""

    def __init__(self, unet, vae_decoder, scheduler, comm: BandComm, guidance_scale: float = 7.5,
                 use_cuda_graph: Optional[bool] = None):
        import os

        from .pipeline import B200Pipeline
        self.comm = comm
        self.pp_unet = PatchParallelUNet(unet, comm)
        self.unet, self.scheduler, self.guidance_scale = unet, scheduler, guidance_scale
        self._single = B200Pipeline(unet, vae_decoder, scheduler, use_cuda_graph=False)
        self.use_cuda_graph = (os.environ.get("B200SD_PP_GRAPH", "1") != "0") if use_cuda_graph is None else use_cuda_graph
        self.step_ms: List[float] = []
        self._key, self._st, self._graph = None, None, None

    def release(self) -> None:
        """drops the captured step.  Call it before the process group is destroyed: a CUDA graph that holds captured
        NCCL kernels keeps their communicators busy, and tearing the group down underneath it does not return."""
        self._graph, self._st, self._key = None, None, None
        torch.cuda.synchronize()

    def _step_body(self, st) -> None:
        sch = self.scheduler
        x_in = st["x"]
        if sch.kind == "euler":
            x_in = st["x"] * st["in_scale"].index_select(0, st["step_idx"].long()).view(1, 1, 1, 1)
        eps2 = self.pp_unet.forward_band(x_in, 2 * st["B"], st["t"], st["ctx_kv"], st["ctx_len"])
        sch.step_cfg(eps2, st["step_idx"], st["x"], st["x0_prev"], self.guidance_scale)

    @torch.no_grad()
    def denoise(self, latents_nhwc: torch.Tensor, ctx2: torch.Tensor, steps: int, time_steps: bool = False) -> torch.Tensor:
        """latents_nhwc fp32 [B, h, w, 4] (the same on every rank), ctx2 [2B, L, D] -> final latents [B, h, w, 4] on every rank"""
        dev, sch, comm = self.unet.device, self.scheduler, self.comm
        B, h, w, _ = latents_nhwc.shape
        r0, rows = comm.band(h)
        sch.set_timesteps(steps, dev)
        coef = sch.coef_table(self.guidance_scale)
        key = (B, h, w, steps, sch.kind, tuple(ctx2.shape), coef.data_ptr())
        if self._key != key:
            self._st = {"B": B, "ctx_len": ctx2.shape[1],
                        "x": torch.zeros(B, rows, w, 4, device=dev, dtype=torch.float32),
                        "x0_prev": torch.zeros(B, rows, w, 4, device=dev, dtype=torch.float32) if sch.needs_x0_prev else None,
                        "t": torch.zeros(1, device=dev, dtype=torch.float32),
                        "step_idx": torch.zeros(1, device=dev, dtype=torch.int32)}
            self._key, self._graph = key, None
        st = self._st
        st["x"].copy_(latents_nhwc[:, r0:r0 + rows].to(dev, torch.float32))
        if sch.init_noise_sigma != 1.0:
            st["x"].mul_(sch.init_noise_sigma)
        if st["x0_prev"] is not None:
            st["x0_prev"].zero_()
        kv = self.unet.project_context(ctx2.to(dev))
        if "ctx_kv" in st:
            for dst, src in zip(st["ctx_kv"], kv):
                dst.copy_(src)
        else:
            st["ctx_kv"] = kv
        t_table = sch.timesteps.to(torch.float32)
        idx_table = torch.arange(steps, device=dev, dtype=torch.int32)
        st["in_scale"] = torch.tensor(sch._in_scale, dtype=torch.float32, device=dev)
        if self.use_cuda_graph and self._graph is None:
            # one eager step on a side stream (lazy initialisation, NCCL communicators for every peer), then the capture;
            # every rank does the same, in the same order
            st["t"].copy_(t_table[0:1]); st["step_idx"].copy_(idx_table[0:1])
            keep = st["x"].clone()
            s = torch.cuda.Stream(device=dev)
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                self._step_body(st)
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=s, capture_error_mode="thread_local"):
                self._step_body(st)
            self._graph = graph
            st["x"].copy_(keep)
            if st["x0_prev"] is not None:
                st["x0_prev"].zero_()
        self.step_ms = []
        for i in range(steps):
            if time_steps:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            st["t"].copy_(t_table[i:i + 1], non_blocking=True)
            st["step_idx"].copy_(idx_table[i:i + 1], non_blocking=True)
            if self._graph is not None:
                self._graph.replay()
            else:
                self._step_body(st)
            if time_steps:
                e1.record()
                e1.synchronize()
                self.step_ms.append(e0.elapsed_time(e1))
        return comm.all_gather_bands(st["x"])

    @torch.no_grad()
    def __call__(self, prompt_embeds: torch.Tensor, negative_prompt_embeds: torch.Tensor, latents: torch.Tensor,
                 num_inference_steps: int = 25, output_type: str = "uint8", time_steps: bool = False):
        from types import SimpleNamespace
        dev = self.unet.device
        lat = latents.to(dev, torch.float32).permute(0, 2, 3, 1).contiguous()
        ctx2 = torch.cat([negative_prompt_embeds.to(dev), prompt_embeds.to(dev)]).to(self.unet.K.op_dtype)
        final = self.denoise(lat, ctx2, num_inference_steps, time_steps)
        if output_type == "latent":
            return SimpleNamespace(images=final.permute(0, 3, 1, 2).clone())
        return SimpleNamespace(images=self._single.decode(final, "uint8"))

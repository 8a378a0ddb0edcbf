"""B200Pipeline -- the denoising loop of diffusers' StableDiffusionPipeline.__call__ (reached by the reference at
app/run-sd.py:140 `pipe(**model_args).images[0]`), with the same keywords and `.images` result, driving
B200UNet / B200Scheduler / B200VAEDecoder.  When real diffusers is importable the three modules can instead be
assigned onto a stock pipeline (see `install_into`), which is the reference's own plugin seam
(app/run-sd.py:108-134).

Per denoising step the device runs ONE CUDA graph: UNet forward (CFG batch 2B, NHWC) + CFG/scheduler kernel.
Timestep and step index live in device memory, so the same graph is replayed for every step.
"""
from __future__ import annotations

import hashlib
from types import SimpleNamespace
from typing import List, Optional, Sequence, Union

import torch

from . import ops
from .scheduler import B200Scheduler
from .unet import B200UNet
from .vae import B200PostQuantConv, B200VAEDecoder


class StandInTextEncoder:
    """PLACEHOLDER used only when no text encoder is given: a prompt is mapped to a deterministic pseudo-embedding
    seeded by its SHA-256.  The real path is `b200sd.clip.PromptEncoder(B200CLIPTextEncoder(...))` (pass it as
    `text_encoder=`); callers with ready embeddings pass `prompt_embeds` / `negative_prompt_embeds`."""

    def __init__(self, dim: int = 1024, length: int = 77):
        self.dim, self.length = dim, length
        self._cache = {}

    def __call__(self, prompts: Sequence[str]) -> torch.Tensor:
        out = []
        for p in prompts:
            if p not in self._cache:
                seed = int.from_bytes(hashlib.sha256(p.encode("utf-8")).digest()[:8], "little") % (2 ** 63)
                g = torch.Generator("cpu").manual_seed(seed)
                self._cache[p] = torch.randn(self.length, self.dim, generator=g)
            out.append(self._cache[p])
        return torch.stack(out)


class B200Pipeline:
    def __init__(self, unet: B200UNet, vae_decoder: B200VAEDecoder, scheduler: Optional[B200Scheduler] = None,
                 text_encoder=None, guidance_scale: float = 7.5, use_cuda_graph: bool = True):
        self.unet, self.vae_decoder = unet, vae_decoder
        self.device = unet.device
        self.scheduler = scheduler or B200Scheduler("ddim", unet.config.prediction_type, self.device)
        self.text_encoder = text_encoder or StandInTextEncoder(unet.config.cross_attention_dim)
        self.guidance_scale = guidance_scale
        self.use_cuda_graph = use_cuda_graph
        self.vae_scale_factor = 8
        self._graph_key = None
        self._graph = None
        self._static = None
        # diffusers-style handle so `pipe.vae.decoder` / `pipe.vae.post_quant_conv` / `pipe.vae.config` resolve
        self.vae = SimpleNamespace(decoder=vae_decoder, post_quant_conv=B200PostQuantConv(),
                                   config=vae_decoder.config, dtype=vae_decoder.dtype)

    def enable_attention_slicing(self, *_a, **_k):
        self.unet.set_attention_slice("auto")

    # ---- one denoising step, graph-captured -------------------------------------------------------------------
    def _step_body(self, st) -> None:
        x_in = st["x"]
        if self.scheduler.kind == "euler":
            x_in = st["x"] * st["in_scale"].index_select(0, st["step_idx"].long()).view(1, 1, 1, 1)
        eps2 = self.unet.forward_nhwc(x_in, 2 * st["B"], st["t"], st["ctx_kv"], st["ctx_len"])
        self.scheduler.step_cfg(eps2, st["step_idx"], st["x"], st["x0_prev"], st["g"])

    def _prepare(self, B: int, h: int, w: int, steps: int, g: float, ctx2: torch.Tensor):
        dev = self.device
        sch = self.scheduler
        sch.set_timesteps(steps, dev)
        coef = sch.coef_table(g)
        # the captured graph holds the addresses of the static buffers and of the scheduler's coefficient table
        key = (B, h, w, steps, g, sch.kind, tuple(ctx2.shape), coef.data_ptr())
        if self._graph_key != key:
            st = {"B": B, "g": g, "ctx_len": ctx2.shape[1],
                  "x": torch.zeros(B, h, w, 4, device=dev, dtype=torch.float32),
                  "x0_prev": torch.zeros(B, h, w, 4, device=dev, dtype=torch.float32) if sch.needs_x0_prev else None,
                  "t": torch.zeros(1, device=dev, dtype=torch.float32),
                  "step_idx": torch.zeros(1, device=dev, dtype=torch.int32),
                  "ctx2": torch.zeros(ctx2.shape, device=dev, dtype=self.unet.K.op_dtype)}
            self._static, self._graph, self._graph_key = st, None, key
        st = self._static
        st["t_table"] = sch.timesteps.to(torch.float32)
        st["idx_table"] = torch.arange(steps, device=dev, dtype=torch.int32)
        scale = torch.tensor(sch._in_scale, dtype=torch.float32)
        if "in_scale" in st and st["in_scale"].shape == scale.shape:
            st["in_scale"].copy_(scale)
        else:
            st["in_scale"] = scale.to(dev)
        st["ctx2"].copy_(ctx2, non_blocking=True)
        # step-invariant cross-attention K/V, recomputed per prompt batch into static buffers
        kv = self.unet.project_context(st["ctx2"])
        if "ctx_kv" in st:
            for dst, src in zip(st["ctx_kv"], kv):
                dst.copy_(src)
        else:
            st["ctx_kv"] = kv
        return st

    @torch.no_grad()
    def denoise(self, latents_nhwc: torch.Tensor, ctx2: torch.Tensor, steps: int,
                guidance_scale: Optional[float] = None) -> torch.Tensor:
        """latents_nhwc: device fp32 [B,h,w,4] unit-variance noise; ctx2: device [2B,L,D] (uncond first).
        Returns the final latents (fp32 NHWC), a view of a static buffer valid until the next call."""
        g = self.guidance_scale if guidance_scale is None else guidance_scale
        B, h, w, _ = latents_nhwc.shape
        st = self._prepare(B, h, w, steps, g, ctx2)
        st["x"].copy_(latents_nhwc)
        if self.scheduler.init_noise_sigma != 1.0:
            st["x"].mul_(self.scheduler.init_noise_sigma)
        if st["x0_prev"] is not None:
            st["x0_prev"].zero_()
        if self.use_cuda_graph and self._graph is None:
            # warm up on a side stream (lazy module init, smem attributes), then capture
            st["t"].copy_(st["t_table"][0:1]); st["step_idx"].copy_(st["idx_table"][0:1])
            keep = st["x"].clone()
            s = torch.cuda.Stream(device=self.device)
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                self._step_body(st)
            torch.cuda.current_stream().wait_stream(s)
            graph = torch.cuda.CUDAGraph()
            # capture on a stream of THIS pipeline's device: torch.cuda.graph's default capture stream is created once per
            # process, on whichever device was current then, and entering it would switch a second replica to that device
            # (thread_local: replicas of other GPUs run in other threads of this process and may allocate or synchronise
            # while this one captures)
            with torch.cuda.graph(graph, stream=s, capture_error_mode="thread_local"):
                self._step_body(st)
            self._graph = graph
            st["x"].copy_(keep)
            if st["x0_prev"] is not None:
                st["x0_prev"].zero_()
        for i in range(steps):
            st["t"].copy_(st["t_table"][i:i + 1], non_blocking=True)
            st["step_idx"].copy_(st["idx_table"][i:i + 1], non_blocking=True)
            if self._graph is not None:
                self._graph.replay()
            else:
                self._step_body(st)
        return st["x"]

    @torch.no_grad()
    def decode(self, latents_nhwc: torch.Tensor, output_type: str = "uint8") -> torch.Tensor:
        img = self.vae_decoder.decode_batched_nhwc(latents_nhwc / self.vae_decoder.config.scaling_factor)
        if output_type == "float":
            return img
        out = torch.empty(img.shape, device=self.device, dtype=torch.uint8)
        ops.denorm_u8(img, out)
        return out

    # ---- embeddings ---------------------------------------------------------------------------------------------
    def encode_prompt(self, prompt: Union[str, List[str]], negative_prompt: Union[None, str, List[str]] = None):
        prompts = [prompt] if isinstance(prompt, str) else list(prompt)
        neg = [""] * len(prompts) if negative_prompt is None else (
            [negative_prompt] * len(prompts) if isinstance(negative_prompt, str) else list(negative_prompt))
        return self.text_encoder(prompts), self.text_encoder(neg)

    # ---- diffusers-compatible call --------------------------------------------------------------------------------
    @torch.no_grad()
    def __call__(self, prompt: Union[None, str, List[str]] = None, height: Optional[int] = None,
                 width: Optional[int] = None, num_inference_steps: int = 50, guidance_scale: Optional[float] = None,
                 negative_prompt=None, num_images_per_prompt: int = 1, eta: float = 0.0, generator=None,
                 latents: Optional[torch.Tensor] = None, prompt_embeds: Optional[torch.Tensor] = None,
                 negative_prompt_embeds: Optional[torch.Tensor] = None, output_type: str = "pil",
                 return_dict: bool = True, **_unused):
        cfg = self.unet.config
        height = height or cfg.sample_size * self.vae_scale_factor
        width = width or cfg.sample_size * self.vae_scale_factor
        if prompt_embeds is None:
            if prompt is None:
                raise ValueError("either `prompt` or `prompt_embeds` is required")
            prompt_embeds, neg = self.encode_prompt(prompt, negative_prompt)
            if negative_prompt_embeds is None:
                negative_prompt_embeds = neg
        elif negative_prompt_embeds is None:
            negative_prompt_embeds = self.text_encoder([""] * prompt_embeds.shape[0])
        if num_images_per_prompt > 1:
            prompt_embeds = prompt_embeds.repeat_interleave(num_images_per_prompt, 0)
            negative_prompt_embeds = negative_prompt_embeds.repeat_interleave(num_images_per_prompt, 0)
        B = prompt_embeds.shape[0]
        h, w = height // 8, width // 8
        if latents is None:
            gen = generator if isinstance(generator, torch.Generator) else None
            if gen is not None and gen.device.type != "cpu":
                latents = torch.randn(B, 4, h, w, generator=gen, device=gen.device)
            else:
                latents = torch.randn(B, 4, h, w, generator=gen)
        dev = self.device
        lat = latents.to(dev, torch.float32, non_blocking=True).permute(0, 2, 3, 1).contiguous()
        ctx2 = torch.cat([negative_prompt_embeds.to(dev, non_blocking=True),
                          prompt_embeds.to(dev, non_blocking=True)]).to(self.unet.K.op_dtype)
        final = self.denoise(lat, ctx2, num_inference_steps, guidance_scale)
        if output_type == "latent":
            images = final.permute(0, 3, 1, 2).clone()
        else:
            u8 = self.decode(final, "uint8")
            if output_type in ("uint8", "pt_uint8"):
                images = u8
            else:
                arr = u8.cpu().numpy()
                if output_type == "pil":
                    from PIL import Image
                    images = [Image.fromarray(a) for a in arr]
                else:  # "np": float32 in [0,1] like diffusers
                    images = arr.astype("float32") / 255.0
        if not return_dict:
            return (images, None)
        return SimpleNamespace(images=images, nsfw_content_detected=None)


def install_into(pipe, unet_sd=None, vae_sd=None, scheduler_kind: str = "ddim", device="cuda",
                 swap_text_encoder: bool = True):
    """The reference's plugin seam (app/run-sd.py:108-134) applied to a stock diffusers pipeline:
    replaces pipe.unet, pipe.text_encoder, pipe.vae.decoder, pipe.vae.post_quant_conv and pipe.scheduler by B200
    modules built from the pipeline's own weights (the same five attributes the reference reassigns)."""
    unet_sd = unet_sd or pipe.unet.state_dict()
    vae_sd = vae_sd or pipe.vae.state_dict()
    ucfg = pipe.unet.config
    pred = getattr(pipe.scheduler.config, "prediction_type", "v_prediction")
    from .config import UNetConfig, VAEConfig
    ucfg2 = UNetConfig.from_any(ucfg)
    ucfg2.prediction_type = pred
    pipe.unet = B200UNet(unet_sd, ucfg2, device)
    pipe.vae.decoder = B200VAEDecoder(vae_sd, VAEConfig.from_any(pipe.vae.config), device)
    pipe.vae.post_quant_conv = B200PostQuantConv()
    pipe.scheduler = B200Scheduler(scheduler_kind, pred, device)
    if swap_text_encoder and getattr(pipe, "text_encoder", None) is not None:
        from .clip import B200CLIPTextEncoder, CLIPTextConfig
        te = pipe.text_encoder
        pipe.text_encoder = B200CLIPTextEncoder(te.state_dict(), CLIPTextConfig.from_any(te.config), device,
                                                out_dtype=getattr(te, "dtype", None))
    return pipe

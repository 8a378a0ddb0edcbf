"""Serving host for the SD2.1 path: the reference's FastAPI surface (app/run-sd.py:153-203) and environment
contract (app/run-sd.py:15-21) on top of B200Pipeline, one engine per GPU.

Kept from the reference: env names APP, POD_NAME, NODEPOOL, MODEL_ID, DEVICE, COMPILED_MODEL_ID, NUM_OF_RUNS_INF;
routes GET /, GET /load/{n_runs}/infer/{n_inf}, POST /genimage, GET /health, GET /readiness, gradio at /serve (only
if gradio is importable); response shapes; the three metric names `<APP>-counter`, `<NODEPOOL>`, `<APP>-latency`
(app/run-sd.py:166-173); the nearest-rank percentile rule of LatencyCollector (app/run-sd.py:84-102).
Added: DEVICE=b200, POST /generate {prompt, num_inference_steps[, height, width]}, a lock around the engine (the
reference shares one stateful pipeline between threadpool handlers without one, app/run-sd.py:137-142).
"""
from __future__ import annotations

import base64
import io
import json
import math
import os
import threading
import time
from typing import Callable, Dict, List, Optional

ENV_KEYS = ("APP", "POD_NAME", "NODEPOOL", "MODEL_ID", "DEVICE", "COMPILED_MODEL_ID", "NUM_OF_RUNS_INF")
CW_NAMESPACE = "hw-agnostic-infer"
WARMUP_PROMPT = "portrait photo of a old warrior chief"
LOAD_PROMPT = "a photo of an astronaut riding a horse on mars"


def read_env(env=os.environ) -> Dict[str, str]:
    """All seven variables are required, as in the reference (KeyError at import if one is unset)."""
    cfg = {k: env[k] for k in ENV_KEYS}
    cfg["NUM_OF_RUNS_INF"] = int(cfg["NUM_OF_RUNS_INF"])
    return cfg


# ---------------------------------------------------------------------------------------------------- metrics
class MetricSink:
    """`put_metric_data` stand-in: boto3/CloudWatch when importable and enabled, else JSON lines on stdout."""

    def __init__(self, namespace: str = CW_NAMESPACE, region: str = "us-west-2", use_cloudwatch: Optional[bool] = None):
        self.namespace = namespace
        self.records: List[dict] = []
        self._cw = None
        want = os.environ.get("B200SD_CLOUDWATCH", "0") == "1" if use_cloudwatch is None else use_cloudwatch
        if want:
            try:
                import boto3  # type: ignore
                self._cw = boto3.client("cloudwatch", region_name=region)
            except Exception:  # boto3 is absent on this box
                self._cw = None

    def publish(self, name: str, value: float, unit: str) -> dict:
        rec = {"Namespace": self.namespace, "MetricName": name, "Value": value, "Unit": unit}
        self.records.append(rec)
        if self._cw is not None:
            return self._cw.put_metric_data(Namespace=self.namespace,
                                            MetricData=[{"MetricName": name, "Value": value, "Unit": unit}])
        print(json.dumps({"metric": rec}), flush=True)
        return rec


# ---------------------------------------------------------------------------------------------------- latency
class LatencyCollector:
    """Wall-clock samples with the reference's percentile rule: position p% of the way into the sorted list,
    rounded up only when the fractional part exceeds one half (app/run-sd.py:94-102)."""

    def __init__(self):
        self.start: Optional[float] = None
        self.latency_list: List[float] = []

    def pre_hook(self, *_):
        self.start = time.time()

    def hook(self, *_):
        self.latency_list.append(time.time() - self.start)

    def percentile(self, percent: float) -> float:
        ordered = sorted(self.latency_list)
        last = len(ordered) - 1
        pos = len(ordered) * percent / 100
        lo, hi = min(math.floor(pos), last), min(math.ceil(pos), last)
        return ordered[hi] if pos - lo > 0.5 else ordered[lo]


def benchmark(n_runs: int, test_name: str, model: Callable, model_inputs: dict, pod_name: str = "") -> str:
    """n_runs timed calls, no warm-up; report string in the reference's format (app/run-sd.py:49-82)."""
    lc = LatencyCollector()
    for _ in range(n_runs):
        lc.pre_hook()
        model(**model_inputs)
        lc.hook()
    parts = [f"Latency P{p}={lc.percentile(p) * 1000:.1f}" for p in (0, 50, 90, 95, 99, 100)]
    report = f"RESULT FOR {test_name} on {pod_name}: " + " ".join(parts)
    print(report, flush=True)
    return report


def serialize_image(image) -> str:
    """PIL image -> base64 PNG (app/run-sd.py:177-181)."""
    buf = io.BytesIO()
    image.save(buf, format="PNG")
    return base64.b64encode(buf.getvalue()).decode("utf-8")


# ---------------------------------------------------------------------------------------------------- engine
class Engine:
    """One pipeline on one GPU behind a lock; `text2img` mirrors app/run-sd.py:137-142."""

    def __init__(self, pipe, default_steps: int):
        self.pipe, self.default_steps = pipe, default_steps
        self._lock = threading.Lock()

    def _on_device(self):
        """Context that makes the pipeline's GPU the current CUDA device of the calling thread.  libb200sd launches on
        the current device and torch's current stream is per device and per thread, so every thread that drives a
        replica (FastAPI threadpool workers, the batcher thread, the warm-up call) must enter this first."""
        import contextlib
        dev = getattr(self.pipe, "device", None)
        if dev is None or getattr(dev, "type", None) != "cuda":
            return contextlib.nullcontext()
        import torch
        return torch.cuda.device(dev)

    def __call__(self, **model_args):
        with self._lock, self._on_device():
            return self.pipe(**model_args)

    def text2img(self, prompt: str, steps: Optional[int] = None):
        t0 = time.time()
        image = self(prompt=prompt, num_inference_steps=steps or self.default_steps).images[0]
        return image, str(time.time() - t0)


# ---------------------------------------------------------------------------------------------------- batching
class _Pending:
    __slots__ = ("prompt", "steps", "event", "image", "error", "t_enqueue", "t_done")

    def __init__(self, prompt: str, steps: int):
        self.prompt, self.steps = prompt, steps
        self.event = threading.Event()
        self.image = self.error = None
        self.t_enqueue, self.t_done = time.time(), None


class BatchingEngine(Engine):
    """dynamic batching in front of one GPU + PNG work off the GPU thread.

    The reference serves one prompt per `pipe(...)` call from sync FastAPI handlers that share a stateful scheduler
    without a lock (app/run-sd.py:137-142, 184-193).  Here handlers enqueue; one worker thread per GPU drains up to
    `max_batch` requests that ask for the same step count (waiting at most `max_wait_ms` for the batch to fill), runs
    ONE pipeline call with `output_type="uint8"` (K7 on the device, 4x less D2H than fp32) and hands each uint8 image
    to a thread pool that builds the PIL image / PNG, so encoding overlaps the next batch on the GPU."""

    def __init__(self, pipe, default_steps: int, max_batch: int = 16, max_wait_ms: float = 5.0, png_workers: int = 4,
                 seed_from_prompt: Optional[bool] = None):
        super().__init__(pipe, default_steps)
        # reproducible serving: the initial noise of a request is drawn from a generator seeded by its prompt, so the same
        # prompt gives the same image whichever replica, batch or batch slot serves it (B200SD_SEED_FROM_PROMPT=1)
        self.seed_from_prompt = (os.environ.get("B200SD_SEED_FROM_PROMPT", "0") == "1") if seed_from_prompt is None \
            else seed_from_prompt
        from concurrent.futures import ThreadPoolExecutor
        import queue
        self.max_batch, self.max_wait_s = max_batch, max_wait_ms / 1e3
        self._q: "queue.Queue[_Pending]" = queue.Queue()
        self._pool = ThreadPoolExecutor(max_workers=png_workers, thread_name_prefix="b200sd-png")
        self.batches: List[int] = []  # sizes of the batches run so far (observability / tests)
        self._pinned = None
        self._stop = False
        self._worker = threading.Thread(target=self._run, name="b200sd-batcher", daemon=True)
        self._worker.start()

    @property
    def queue_depth(self) -> int:
        return self._q.qsize()

    def close(self):
        self._stop = True
        self._q.put(None)  # type: ignore[arg-type]
        self._worker.join(timeout=5)
        self._pool.shutdown(wait=False)

    def _take_batch(self) -> List[_Pending]:
        import queue
        first = self._q.get()
        if first is None:
            return []
        batch, deadline, later = [first], time.time() + self.max_wait_s, []
        while len(batch) < self.max_batch:
            try:
                nxt = self._q.get(timeout=max(0.0, deadline - time.time()))
            except queue.Empty:
                break
            if nxt is None:
                self._stop = True
                break
            (batch if nxt.steps == first.steps else later).append(nxt)
        for r in later:  # different step count: next batch
            self._q.put(r)
        return batch

    def _run(self):
        while not self._stop:
            batch = self._take_batch()
            if not batch:
                continue
            self.batches.append(len(batch))
            try:
                extra = {"latents": self._prompt_latents([r.prompt for r in batch])} if self.seed_from_prompt else {}
                with self._lock, self._on_device():
                    out = self.pipe(prompt=[r.prompt for r in batch], num_inference_steps=batch[0].steps,
                                    output_type="uint8", **extra).images
                    arr = self._to_host(out)
                for r, a in zip(batch, arr):
                    self._pool.submit(self._finish, r, a)
            except Exception as e:  # every waiter of the batch sees the failure (HTTP 500 in the handler)
                for r in batch:
                    r.error = e
                    r.event.set()

    def _prompt_latents(self, prompts: List[str]):
        import hashlib

        import torch
        hw = self.pipe.unet.config.sample_size
        out = []
        for p in prompts:
            seed = int.from_bytes(hashlib.sha256(p.encode("utf-8")).digest()[:8], "little") % (2 ** 63)
            out.append(torch.randn(4, hw, hw, generator=torch.Generator("cpu").manual_seed(seed)))
        return torch.stack(out)

    def _to_host(self, out):
        """device uint8 images -> numpy through a pinned staging buffer (async copy + one stream sync, no pageable
        bounce); fake pipelines in the CPU tests hand back arrays already"""
        if not hasattr(out, "cpu"):
            return out
        if not out.is_cuda:
            return out.numpy()
        import torch
        if self._pinned is None or self._pinned.shape[1:] != out.shape[1:] or self._pinned.shape[0] < out.shape[0]:
            self._pinned = torch.empty((max(out.shape[0], self.max_batch),) + tuple(out.shape[1:]), dtype=out.dtype,
                                       pin_memory=True)
        host = self._pinned[: out.shape[0]]
        host.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return host.numpy().copy()  # the staging buffer is reused by the next batch while PNG workers still read this one

    @staticmethod
    def _finish(r: _Pending, arr):
        try:
            from PIL import Image
            r.image = Image.fromarray(arr)
        except Exception as e:
            r.error = e
        r.t_done = time.time()
        r.event.set()

    def submit(self, prompt: str, steps: Optional[int] = None) -> _Pending:
        r = _Pending(prompt, steps or self.default_steps)
        self._q.put(r)
        return r

    def text2img(self, prompt: str, steps: Optional[int] = None):
        r = self.submit(prompt, steps)
        r.event.wait()
        if r.error is not None:
            raise r.error
        return r.image, str(r.t_done - r.t_enqueue)


class Dispatcher:
    """One engine per GPU of the box, requests go to the engine with the shortest queue (ties: round-robin): the in-box
    analogue of the reference's N one-GPU pods behind ALB round-robin (sd21-equal-routing-ing.yaml:33).  No collective."""

    def __init__(self, engines: List[Engine]):
        if not engines:
            raise ValueError("Dispatcher needs at least one engine")
        self.engines, self.default_steps = list(engines), engines[0].default_steps
        self._rr, self._lock = 0, threading.Lock()
        self.routed = [0] * len(engines)

    def pick(self) -> int:
        with self._lock:
            n = len(self.engines)
            order = [(self._rr + i) % n for i in range(n)]
            best = min(order, key=lambda i: getattr(self.engines[i], "queue_depth", 0))
            self._rr = (best + 1) % n
            self.routed[best] += 1
            return best

    def text2img(self, prompt: str, steps: Optional[int] = None):
        return self.engines[self.pick()].text2img(prompt, steps)

    def __call__(self, **model_args):
        return self.engines[self.pick()](**model_args)


def _random_clip_state_dict(ccfg, seed: int):
    """seeded text-tower weights in transformers' CLIPTextModel layout"""
    import torch
    g = torch.Generator("cpu").manual_seed(seed)
    c, f = ccfg.hidden_size, ccfg.intermediate_size
    csd = {"text_model.embeddings.token_embedding.weight": torch.randn(ccfg.vocab_size, c, generator=g),
           "text_model.embeddings.position_embedding.weight": 0.1 * torch.randn(ccfg.max_position_embeddings, c, generator=g),
           "text_model.final_layer_norm.weight": torch.ones(c), "text_model.final_layer_norm.bias": torch.zeros(c)}
    for i in range(ccfg.num_hidden_layers):
        q = f"text_model.encoder.layers.{i}."
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            csd[q + f"self_attn.{n}.weight"] = torch.randn(c, c, generator=g) * c ** -0.5
            csd[q + f"self_attn.{n}.bias"] = torch.zeros(c)
        for n in ("layer_norm1", "layer_norm2"):
            csd[q + n + ".weight"], csd[q + n + ".bias"] = torch.ones(c), torch.zeros(c)
        csd[q + "mlp.fc1.weight"], csd[q + "mlp.fc1.bias"] = torch.randn(f, c, generator=g) * c ** -0.5, torch.zeros(f)
        csd[q + "mlp.fc2.weight"], csd[q + "mlp.fc2.bias"] = torch.randn(c, f, generator=g) * f ** -0.5, torch.zeros(c)
    return csd


def build_random_init_pipeline(model_id: str, device: str = "cuda:0", seed: int = 0):
    """There are no checkpoints on this box (no network): builds the SD2.1 architecture named by `model_id`
    (`...-base` = 512^2 epsilon-prediction, otherwise 768^2 v-prediction) with seeded random weights.
    With real weights, pass a diffusers-layout state_dict to B200UNet / B200VAEDecoder instead."""
    import torch
    import torch.nn as nn

    from . import B200Pipeline, B200Scheduler, B200UNet, B200VAEDecoder, UNetConfig, VAEConfig
    from ._randinit import random_state_dicts

    base = model_id.rstrip("/").endswith("-base")
    ucfg = UNetConfig(sample_size=64 if base else 96, prediction_type="epsilon" if base else "v_prediction")
    usd, vsd = random_state_dicts(ucfg, VAEConfig(), seed)
    dev = torch.device(device)
    unet, vae = B200UNet(usd, ucfg, dev), B200VAEDecoder(vsd, VAEConfig(), dev)
    text = None
    if os.environ.get("B200SD_TEXT_ENCODER", "clip") == "clip":
        # the SD2.1 text tower with seeded weights in transformers' layout + the placeholder tokenizer (no vocabulary here)
        from .clip import B200CLIPTextEncoder, CLIPTextConfig, PromptEncoder
        ccfg = CLIPTextConfig()
        csd = _random_clip_state_dict(ccfg, seed + 2)
        text = PromptEncoder(B200CLIPTextEncoder(csd, ccfg, dev))
    return B200Pipeline(unet, vae, B200Scheduler("ddim", ucfg.prediction_type, dev), text_encoder=text)


def build_tiny_pipeline(model_id: str, device: str = "cuda:0", seed: int = 0):
    """B200SD_MODEL_SIZE=tiny: the same modules on a small UNet / VAE / text tower (16x16 latents, 128 x 128 images), so a
    test can bring the whole host up -- warm-up, batcher, routes -- in seconds.  Not a serving configuration."""
    import torch

    from . import B200Pipeline, B200Scheduler, B200UNet, B200VAEDecoder, UNetConfig, VAEConfig
    from ._randinit import random_state_dicts
    from .clip import B200CLIPTextEncoder, CLIPTextConfig, PromptEncoder

    ucfg = UNetConfig(sample_size=16, block_out_channels=(64, 128, 128, 128), attention_head_dim=(1, 2, 2, 2),
                      cross_attention_dim=128)
    vcfg = VAEConfig(block_out_channels=(64, 64, 128, 128))
    usd, vsd = random_state_dicts(ucfg, vcfg, seed)
    dev = torch.device(device)
    ccfg = CLIPTextConfig(hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=2)
    text = PromptEncoder(B200CLIPTextEncoder(_random_clip_state_dict(ccfg, seed + 2), ccfg, dev))
    return B200Pipeline(B200UNet(usd, ucfg, dev), B200VAEDecoder(vsd, vcfg, dev),
                        B200Scheduler("ddim", ucfg.prediction_type, dev), text_encoder=text)


def create_app(engine: Engine, cfg: Dict[str, str], sink: Optional[MetricSink] = None):
    """FastAPI app with the reference's routes and payloads."""
    from fastapi import FastAPI

    sink = sink or MetricSink()
    app = FastAPI()
    model_id, pod, device = cfg["MODEL_ID"], cfg["POD_NAME"], cfg["DEVICE"]

    @app.get("/")
    def read_main():
        return {"message": "This is" + model_id + " pod " + pod + " in AWS EC2 " + device +
                           " instance; try /load/{n_runs}/infer/{n_inf}; /genimage http post with user prompt "}

    @app.get("/load/{n_runs}/infer/{n_inf}")
    def load(n_runs: int, n_inf: int):
        t0 = time.time()
        report = benchmark(n_runs, "stable_diffusion_512", engine,
                           {"prompt": LOAD_PROMPT, "num_inference_steps": n_inf}, pod)
        total = time.time() - t0
        sink.publish(cfg["APP"] + "-counter", 1, "Count")
        sink.publish(cfg["NODEPOOL"], 1, "Count")
        sink.publish(cfg["APP"] + "-latency", total, "Seconds")
        return {"message": "benchmark report:" + report}

    @app.post("/genimage")
    def generate_image_post(request: dict):
        prompt = request.get("prompt")
        image, latency = engine.text2img(prompt)
        return {"prompt": prompt, "response": serialize_image(image), "latency": latency}

    @app.post("/generate")
    def generate(request: dict):
        """superset endpoint in the shape of the reference's Flux model-API (app/flux_model_api.py:331-360)"""
        prompt = request.get("prompt")
        args = {"prompt": prompt, "num_inference_steps": int(request.get("num_inference_steps", engine.default_steps))}
        for k in ("height", "width"):
            if request.get(k) is not None:
                args[k] = int(request[k])
        t0 = time.time()
        image = engine(**args).images[0]
        return {"image": serialize_image(image), "execution_time": time.time() - t0}

    @app.get("/health")
    def healthy():
        return {"message": pod + "is healthy"}

    @app.get("/readiness")
    def ready():
        return {"message": pod + "is ready"}

    try:  # gradio is optional (absent on this box)
        import gradio as gr  # type: ignore

        io_ = gr.Interface(fn=engine.text2img, inputs=["text"], outputs=[gr.Image(height=512, width=512), "text"],
                           title=model_id + " on " + device + "; pod name " + pod)
        app = gr.mount_gradio_app(app, io_, path="/serve")
    except Exception:
        pass
    return app


def create_app_from_env():
    """`uvicorn b200sd.serving:create_app_from_env --factory` -- the analogue of `uvicorn run-sd:app`
    (app/run-sd.sh:14).  Builds the pipeline, runs the reference's warm-up inference, then serves."""
    cfg = read_env()
    if cfg["DEVICE"] not in ("b200", "cuda"):
        raise RuntimeError(f"DEVICE={cfg['DEVICE']!r}: this host serves DEVICE=b200 (sm_100a) only")
    # B200SD_GPUS=N serves N replicas (cuda:0..N-1) behind one Dispatcher; each replica batches up to B200SD_MAX_BATCH
    # concurrent requests into one pipeline call.  Defaults: 1 GPU (B200SD_CUDA_DEVICE), batch 16.
    n_gpus = int(os.environ.get("B200SD_GPUS", "1"))
    devices = [os.environ.get("B200SD_CUDA_DEVICE", "cuda:0")] if n_gpus == 1 else [f"cuda:{i}" for i in range(n_gpus)]
    engines = []
    import torch
    builder = build_tiny_pipeline if os.environ.get("B200SD_MODEL_SIZE", "full") == "tiny" else build_random_init_pipeline
    for d in devices:
        with torch.cuda.device(torch.device(d)):  # weights, workspaces and the lazily set kernel attributes land on GPU d
            pipe = builder(cfg["MODEL_ID"], d)
        pipe.enable_attention_slicing()
        eng = BatchingEngine(pipe, cfg["NUM_OF_RUNS_INF"], max_batch=int(os.environ.get("B200SD_MAX_BATCH", "16")),
                             max_wait_ms=float(os.environ.get("B200SD_BATCH_WAIT_MS", "5")))
        eng(prompt=WARMUP_PROMPT, num_inference_steps=cfg["NUM_OF_RUNS_INF"])  # readiness only after warm-up
        engines.append(eng)
    return create_app(engines[0] if len(engines) == 1 else Dispatcher(engines), cfg)


# ---------------------------------------------------------------------------------------------------- sharding
def shard_range(n_items: int, rank: int, world: int) -> range:
    """Data-parallel serving: independent prompts are split contiguously across replicas (one per GPU), the
    in-box analogue of the reference's N pods behind ALB round-robin (sd21-equal-routing-ing.yaml:33).
    No collective on the data path."""
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))

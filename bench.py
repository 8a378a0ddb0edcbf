#!/usr/bin/env python
"""bench.py -- SD2.1 768^2 / 25-step images per second on N B200s (BASELINE.json metric), one JSON line on rank 0.

A "step" is one pass of the hot path over one batch: 25 denoising steps (UNet at CFG batch 2B + CFG/scheduler
kernel, replayed as one CUDA graph per step) + VAE decode + uint8 post-process.
  value : whole-job images/s with latents and text embeddings already resident in HBM (CUDA-event timed)
  e2e   : the same through the public API  pipe(prompt_embeds=, negative_prompt_embeds=, latents=, ...)  with
          pinned HOST inputs and the uint8 images read back to the host inside the timed region
  roofline : dominant kernel class, algorithmic FLOPs / CUDA-event time per launch, vs the measured bf16 peak
  cpu_baseline : the fp32 oracle (a port: diffusers is not installable here) on this box's host cores, bounded sample
`--impl reference` times that CPU oracle alone (the reference ships no other implementation of the path).

N > 1 (torchrun): one replica per GPU, independent prompts, no data-path collective; value = all images / max time.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG = ROOT / "scalable-hw-agnostic-inference_b200"
for p in (str(ROOT), str(PKG)):
    if p not in sys.path:
        sys.path.insert(0, p)

UNET_TFLOP = {64: 1.6085, 96: 4.2982, 128: 9.3500}    # per image-step, CFG batch 2
VAE_TFLOP = {64: 2.5145, 96: 5.7543, 128: 10.4704}    # per image


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5, help="timed passes of the hot path")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=16,
                    help="images per GPU per pass: 16 = BASELINE configs[2], the throughput-serving config the images/s "
                         "metric at 1/2/4/8 GPUs is quoted on; 1 = configs[1], single-request latency")
    ap.add_argument("--denoise-steps", type=int, default=25)
    ap.add_argument("--size", type=int, default=768)
    ap.add_argument("--scheduler", default="ddim", choices=["ddim", "dpm++2m", "euler"])
    ap.add_argument("--operand-dtype", default=None, choices=[None, "fp16", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-comparator", action="store_true",
                    help="skip the same-box stock-torch bf16 comparator (tests/torch_gpu_baseline.py) and the bf16-operand sub-line")
    ap.add_argument("--comparator-compile", default="off", choices=["on", "off"],
                    help="torch.compile arm of the comparator (the reference's DEVICE=triton): max-autotune of the UNet takes "
                         "10+ minutes of GPU time, so it is opt-in; a run with it is kept in profiles/")
    ap.add_argument("--patch-parallel", action="store_true",
                    help="BASELINE configs[4]: ONE request split over the ranks by latent rows (NCCL K/V all-gather, halo "
                         "rows, GroupNorm all-reduce); use with --size 1024 --batch 1 under torchrun.  Not the default metric.")
    ap.add_argument("--profile-only", action="store_true", help="one eager profiled pass, no timing (for ncu runs)")
    ap.add_argument("--shape-profile", action="store_true", help="per-(kernel, shape) CUDA-event times of one eager pass")
    return ap.parse_args()


def peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        d = json.loads(f.read_text())
        return {"tflops": float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1590.0))), "hbm_gbs": float(d.get("hbm_gbs", 6650.0)),
                "source": "measured (MEASURED_PEAKS.json, sustained bf16 GEMM)"}
    return {"tflops": 1590.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """samples nvidia-smi clocks and throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------- CPU oracle leg
def cpu_oracle_sample(size: int, denoise_steps: int, n_unet: int = 1):
    """fp32 oracle on all host cores: n_unet UNet forwards (CFG batch 2) + one VAE decode at `size`, extrapolated to a
    full image.  Returns (images_per_s, info)."""
    import torch
    from oracle import sd21 as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hw = size // 8
    unet = O.seeded_init_(O.UNet2DCondition(O.sd21_unet_config(sample_size=hw)), 0).eval()
    vae = O.seeded_init_(O.VAEDecoder(O.VAEConfig()), 1).eval()
    lat, ctx = O.seeded_inputs(1, hw, 1024)
    x2 = torch.cat([lat, lat])
    t_unet = []
    with torch.no_grad():
        for i in range(n_unet):
            t0 = time.perf_counter()
            unet(x2, 961.0, ctx)
            t_unet.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        vae.decode(lat / 0.18215)
        t_vae = time.perf_counter() - t0
    tu = min(t_unet)
    per_image = denoise_steps * tu + t_vae
    return 1.0 / per_image, {"unet_step_s": tu, "vae_decode_s": t_vae, "cores": cores, "unet_times_s": t_unet}


def run_reference(args):
    """`--impl reference`: the reference's own implementation of the path is diffusers on torch, which cannot be
    installed here (no network, not in the wheelhouse); the timed stand-in is the fp32 CPU oracle, all host threads.

    A "step" of this arm is a BOUNDED SAMPLE of the workload: one UNet forward of one image (CFG batch 2) at the
    workload's resolution.  `steps` / `warmup` / `ms_per_step` on the line describe exactly those timed forwards (capped
    by a wall-clock budget), `value` is the images/s they imply for the full workload (denoise_steps forwards + one
    timed VAE decode per image) and is flagged `extrapolated`."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import sd21 as O
    budget_s = float(os.environ.get("B200SD_REF_BUDGET_S", "200"))
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hw = args.size // 8
    t_begin = time.perf_counter()
    unet = O.seeded_init_(O.UNet2DCondition(O.sd21_unet_config(sample_size=hw)), 0).eval()
    vae = O.seeded_init_(O.VAEDecoder(O.VAEConfig()), 1).eval()
    lat, ctx = O.seeded_inputs(1, hw, 1024)
    x2 = torch.cat([lat, lat])
    times, n_warm = [], 0
    with torch.no_grad():
        t0 = time.perf_counter()
        vae.decode(lat / 0.18215)
        t_vae = time.perf_counter() - t0
        n_warm = min(2, max(1, args.warmup))  # at most two untimed forwards: each costs seconds of host time
        for _ in range(n_warm):
            unet(x2, 961.0, ctx)
        for _ in range(max(1, args.steps)):
            t0 = time.perf_counter()
            unet(x2, 961.0, ctx)
            times.append(time.perf_counter() - t0)
            if time.perf_counter() - t_begin + 1.5 * times[-1] > budget_s:
                break
    tu = statistics.mean(times)
    per_image = args.denoise_steps * tu + t_vae
    ips = 1.0 / per_image
    sample = (f"{len(times)} timed UNet forwards of one image (CFG batch 2, mean taken; {n_warm} untimed before) + 1 VAE "
              f"decode at {args.size}^2 fp32; images/s = 1 / ({args.denoise_steps} x forward + decode)")
    line = {"impl": "reference", "metric": "sd21_images_per_sec", "value": ips, "unit": "images/s", "n_gpus": args.gpus,
            "steps": len(times), "warmup": n_warm, "ms_per_step": 1000.0 * tu, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "extrapolated": True,
            "step_is": "one fp32 UNet forward of one image (CFG batch 2) on the host cores -- a bounded sample, not a pass",
            "requested": {"steps": args.steps, "warmup": args.warmup, "budget_s": budget_s},
            # same workload as the B200 arm; the bounded CPU sample runs one image (CFG batch 2) at a time -- the
            # oracle's images/s does not depend on how many prompts are queued
            "config": dict(workload_config(args, args.batch), cpu_sample_batch=1),
            "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample,
                             "unet_step_s": tu, "unet_step_s_min": min(times), "vae_decode_s": t_vae},
            "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    _emit(line)


def workload_config(args, batch):
    return {"workload": f"SD2.1 {args.size}x{args.size}, {args.denoise_steps} {args.scheduler} steps, CFG 7.5, "
                        f"batch {batch} per GPU, random-init weights (BASELINE configs[{1 if batch == 1 else 2}])",
            "size": args.size, "denoise_steps": args.denoise_steps, "batch_per_gpu": batch, "scheduler": args.scheduler,
            "l2": "no explicit flush: each pass streams 1.8 GB of weights plus activations, far beyond the 126 MB L2",
            "parallelism": f"dp{args.gpus} (one replica per GPU, no collective)"}


# ---------------------------------------------------------------------------------------------------- GPU arm
def run_b200(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.operand_dtype:
        os.environ["B200SD_OPERAND_DTYPE"] = args.operand_dtype
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=dev)
        dist = dist_mod

    import b200sd
    from b200sd import ops
    from b200sd._randinit import random_state_dicts

    hw = args.size // 8
    B = args.batch
    ucfg = b200sd.UNetConfig(sample_size=hw, prediction_type="v_prediction" if hw == 96 else "epsilon")
    vcfg = b200sd.VAEConfig()
    usd, vsd = random_state_dicts(ucfg, vcfg, 0)
    unet = b200sd.B200UNet(usd, ucfg, dev)
    vae = b200sd.B200VAEDecoder(vsd, vcfg, dev)
    del usd, vsd
    pipe = b200sd.B200Pipeline(unet, vae, b200sd.B200Scheduler(args.scheduler, ucfg.prediction_type, dev))
    op_name = {torch.float16: "fp16", torch.bfloat16: "bf16"}[unet.K.op_dtype]

    # synthetic inputs: distinct seeded latents / embeddings per rank (independent prompts), pinned on the host
    g = torch.Generator("cpu").manual_seed(1234 + rank)
    lat_host = torch.randn(B, 4, hw, hw, generator=g).pin_memory()
    cond_host = torch.randn(B, 77, 1024, generator=g).pin_memory()
    unc_host = torch.randn(1, 77, 1024, generator=torch.Generator("cpu").manual_seed(7)).expand(B, -1, -1).contiguous().pin_memory()
    lat_dev = lat_host.to(dev).permute(0, 2, 3, 1).contiguous()
    ctx2_dev = torch.cat([unc_host.to(dev), cond_host.to(dev)]).to(unet.K.op_dtype)

    def one_pass_device():
        final = pipe.denoise(lat_dev, ctx2_dev, args.denoise_steps)
        return pipe.decode(final, "uint8")

    def one_pass_e2e():
        res = pipe(prompt_embeds=cond_host, negative_prompt_embeds=unc_host, latents=lat_host,
                   num_inference_steps=args.denoise_steps, height=args.size, width=args.size, output_type="uint8")
        return res.images.cpu()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    if args.profile_only:
        pipe.use_cuda_graph = False
        ops._taglog = []
        one_pass_device()
        torch.cuda.synchronize()
        out = Path(os.environ.get("B200SD_TAGLOG", str(ROOT / "gpurun_out" / "taglog.json")))
        out.parent.mkdir(exist_ok=True)
        out.write_text(json.dumps(ops._taglog))
        return
    if args.shape_profile:
        pipe.use_cuda_graph = False
        one_pass_device()
        ops.profile_begin()
        one_pass_device()
        prof = ops.profile_end(by_shape=True)
        tot = sum(d["ms"] for d in prof.values())
        print(f"total {tot:.2f} ms over {sum(d['launches'] for d in prof.values())} launches")
        for k, d in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])[:60]:
            rate = d["work"] / (d["ms"] * 1e-3) / 1e12
            print(f"{d['ms']:9.3f} ms {100 * d['ms'] / tot:5.1f}%  n={d['launches']:4d}  avg {1000 * d['ms'] / d['launches']:8.1f} us  "
                  f"{rate:8.2f} T(FLOP|B)/s  {k}")
        return

    for _ in range(max(args.warmup, 1)):
        one_pass_device()
    torch.cuda.synchronize()

    # ---- launches per pass: graph replays carry the launches counted at capture
    n0 = ops.launch_count()
    eager_pipe_flag = pipe.use_cuda_graph
    pipe.use_cuda_graph = False
    saved_graph, pipe._graph = pipe._graph, None
    one_pass_device()
    torch.cuda.synchronize()
    launches_per_pass = ops.launch_count() - n0
    pipe.use_cuda_graph, pipe._graph = eager_pipe_flag, saved_graph

    # ---- timed region 1: device-resident inputs
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        one_pass_device()
    ev1.record()
    torch.cuda.synchronize()
    ms_dev = ev0.elapsed_time(ev1)
    clk = clocks.stop()
    barrier()

    # ---- timed region 2: end to end through the public API with host buffers
    for _ in range(2):
        one_pass_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        img = one_pass_e2e()
    torch.cuda.synchronize()
    ms_e2e = (time.perf_counter() - t0) * 1000.0
    barrier()
    h2d = lat_host.numel() * 4 + cond_host.numel() * 4 + unc_host.numel() * 4
    d2h = img.numel()

    if dist is not None:
        t = torch.tensor([ms_dev, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e = float(t[0]), float(t[1])

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- per-kernel-class profile (eager, every launch bracketed by CUDA events on the launching stream)
    pipe.use_cuda_graph = False
    pipe._graph, saved_graph = None, pipe._graph
    ops.profile_begin()
    one_pass_device()
    prof = ops.profile_end()
    pipe.use_cuda_graph, pipe._graph = True, saved_graph
    total_ms = sum(d["ms"] for d in prof.values())
    pk = peaks()
    classes = {}
    for kind, d in sorted(prof.items(), key=lambda kv: -kv[1]["ms"]):
        tensor = kind in ("gemm", "attention")
        rate = d["work"] / (d["ms"] * 1e-3) if d["ms"] > 0 else 0.0
        classes[kind] = {"launches": d["launches"], "ms": round(d["ms"], 3), "share": round(d["ms"] / total_ms, 4),
                         ("tflops" if tensor else "gbs"): round(rate / (1e12 if tensor else 1e9), 2)}
    dom = max(prof.items(), key=lambda kv: kv[1]["ms"])
    dk, dd = dom
    tensor = dk in ("gemm", "attention")
    achieved = dd["work"] / (dd["ms"] * 1e-3) / (1e12 if tensor else 1e9)
    peak = pk["tflops"] if tensor else pk["hbm_gbs"]
    traffic = None
    tf = ROOT / "profiles" / "roofline_traffic.json"
    if tf.exists():
        try:
            # bytes per launch (dram read + write, averaged over every launch of the class in one ncu'd pass)
            traffic = json.loads(tf.read_text()).get(dk)
            if isinstance(traffic, dict):
                traffic = traffic.get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"kernel": {"gemm": "gemm_kernel (implicit-GEMM conv/linear, tcgen05)", "attention": "attention_kernel"}.get(dk, dk),
                "bound": "tensor" if tensor else "hbm", "achieved": round(achieved, 2), "peak": peak,
                "unit": "TFLOP/s" if tensor else "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                "peak_source": pk["source"], "launches_per_pass": dd["launches"],
                "avg_launch_ms": round(dd["ms"] / dd["launches"], 4),
                "how": "sum of algorithmic work over all launches of this kernel in one pass / sum of their CUDA-event "
                       "durations (eager pass, events on the launching stream)",
                "classes": classes}

    images = B * world
    value = images * args.steps / (ms_dev * 1e-3)
    e2e_val = images * args.steps / (ms_e2e * 1e-3)
    tflop_img = args.denoise_steps * UNET_TFLOP.get(hw, 0) + VAE_TFLOP.get(hw, 0)
    line = {"metric": "sd21_images_per_sec", "value": round(value, 4), "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_dev / args.steps, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": op_name, "data": "synthetic",
            "config": workload_config(args, B),
            "unet_step_ms": round((ms_dev / args.steps) * (1 - classes_share_vae(prof, pipe)) / args.denoise_steps, 3),
            "whole_image_tensor_roofline_frac": round(value * tflop_img / (world * pk["tflops"]), 4) if tflop_img else None,
            "e2e": {"value": round(e2e_val, 4), "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": round(ms_e2e / args.steps, 3)},
            "gpu_launches": int(launches_per_pass * args.steps),
            "clocks": clk, "roofline": roofline}
    if B != 1:
        # BASELINE configs[1] beside the headline: one request at batch 1 through the same pipeline (own CUDA graph)
        lat1 = lat_dev[:1].contiguous()
        ctx1 = torch.cat([ctx2_dev[:1], ctx2_dev[B:B + 1]]).contiguous()
        for _ in range(2):
            pipe.decode(pipe.denoise(lat1, ctx1, args.denoise_steps), "uint8")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            pipe.decode(pipe.denoise(lat1, ctx1, args.denoise_steps), "uint8")
        e1.record()
        torch.cuda.synchronize()
        line["single_request"] = {"batch": 1, "ms_per_image": round(e0.elapsed_time(e1) / 3, 3),
                                  "images_per_s": round(3000.0 / e0.elapsed_time(e1), 4)}
        # UNet step alone at batch 1 (the graph-replayed denoising loop without the VAE decode)
        e0.record()
        pipe.denoise(lat1, ctx1, args.denoise_steps)
        e1.record()
        torch.cuda.synchronize()
        line["single_request"]["unet_step_ms"] = round(e0.elapsed_time(e1) / args.denoise_steps, 3)
        # self check: image 0 of the batch-B pass against the same prompt/latent run alone at batch 1 -- other tilings
        # (no CTA pairs, split-K at the low levels), so not bit-equal: the two differ by fp32 summation order and the
        # 16-bit roundings that flips, compounded over the schedule.  PSNR on the uint8 images, peak 255.
        img_b = one_pass_device()[0].float()
        img_1 = pipe.decode(pipe.denoise(lat1, ctx1, args.denoise_steps), "uint8")[0].float()
        mse = float(((img_b - img_1) ** 2).mean())
        psnr = 99.0 if mse == 0 else 10.0 * __import__("math").log10(255.0 ** 2 / mse)
        spread = float(img_b.std())
        ok = psnr >= 40.0 and spread > 1.0   # (a constant image would also "match")
        line["self_check"] = "ok" if ok else "FAILED"
        line["self_check_detail"] = {"what": f"image 0 of the batch-{B} pass vs the same request at batch 1 (uint8 PSNR >= 40 dB)",
                                     "psnr_db": round(psnr, 2), "max_abs_u8": int((img_b - img_1).abs().max()),
                                     "image_std_u8": round(spread, 2)}
    if not args.no_comparator and world == 1:
        # ---- the same throughput with the reference's storage dtype (bf16 operands): driver-observed, not asserted
        usd2, _ = random_state_dicts(ucfg, vcfg, 0)
        unet_bf = b200sd.B200UNet(usd2, ucfg, dev, op_dtype=torch.bfloat16)
        del usd2
        pipe_bf = b200sd.B200Pipeline(unet_bf, vae, b200sd.B200Scheduler(args.scheduler, ucfg.prediction_type, dev))
        ctx_bf = ctx2_dev.to(torch.bfloat16)
        for _ in range(2):
            pipe_bf.decode(pipe_bf.denoise(lat_dev, ctx_bf, args.denoise_steps), "uint8")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            pipe_bf.decode(pipe_bf.denoise(lat_dev, ctx_bf, args.denoise_steps), "uint8")
        e1.record()
        torch.cuda.synchronize()
        line["bf16_operands"] = {"value": round(B * 3000.0 / e0.elapsed_time(e1), 4), "unit": "images/s",
                                 "ms_per_step": round(e0.elapsed_time(e1) / 3, 3), "passes": 3,
                                 "note": "same kernels with bf16 tensor-core operands (the reference's DTYPE, app/run-sd.py:39)"}
        del pipe_bf, unet_bf
    if not args.no_cpu_baseline and world == 1:
        ips, info = cpu_oracle_sample(args.size, args.denoise_steps, n_unet=2)
        line["cpu_baseline"] = {"value": ips, "unit": "images/s", "cores": info["cores"], "kind": "port",
                                "sample": f"2 UNet forwards (CFG batch 2, best taken) + 1 VAE decode at {args.size}^2 fp32, "
                                          f"extrapolated to {args.denoise_steps} steps/image",
                                "unet_step_s": round(info["unet_step_s"], 3), "vae_decode_s": round(info["vae_decode_s"], 3)}
    if not args.no_comparator and world == 1:
        # free this process's GPU memory first: the comparator builds its own fp32 -> bf16 copy of the model
        del pipe, unet, vae
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        line["gpu_comparator"] = gpu_comparator(args, local)
    _emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def gpu_comparator(args, gpu_index: int) -> dict:
    """Same box, same run: stock PyTorch bf16 on the same architecture -- eager (cuDNN / cuBLAS / fused SDPA: the
    reference's DEVICE=cuda, app/run-sd.py:106-107) and torch.compile(mode="max-autotune-no-cudagraphs") of the UNet
    (DEVICE=triton, app/run-sd.py:109-116).  Runs tests/torch_gpu_baseline.py in a subprocess (it is test
    infrastructure: it moves the fp32 oracle's modules to the GPU); a leg that fails or runs out of time reports why."""
    out = {"what": "stock torch bf16 on the oracle's modules, same GPU, same process tree", "unit": "images/s"}
    script = str(ROOT / "tests" / "torch_gpu_baseline.py")
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", str(gpu_index)))
    legs = [("batch16_eager", [str(args.batch)], 150), ("batch1_eager", ["1"], 90)]
    if args.comparator_compile == "on":
        legs += [("batch16_compile", [str(args.batch), "--compile"], 600), ("batch1_compile", ["1", "--compile"], 420)]
    for name, extra, limit in legs:
        t0 = time.perf_counter()
        try:
            r = subprocess.run([sys.executable, script, *extra, "--size", str(args.size), "--steps", str(args.denoise_steps)],
                               capture_output=True, text=True, timeout=limit, env=env)
            last = [ln for ln in r.stdout.strip().splitlines() if ln.startswith("{")]
            out[name] = json.loads(last[-1]) if r.returncode == 0 and last else \
                {"unavailable": f"exit {r.returncode}: {(r.stderr or r.stdout).strip().splitlines()[-1][:200] if (r.stderr or r.stdout).strip() else ''}"}
        except subprocess.TimeoutExpired:
            out[name] = {"unavailable": f"timed out after {limit} s"}
        except Exception as e:  # noqa: BLE001
            out[name] = {"unavailable": repr(e)[:200]}
        out[name]["wall_s"] = round(time.perf_counter() - t0, 1)
    return out


def classes_share_vae(prof, pipe) -> float:
    """fraction of one pass spent outside the denoising loop (VAE decode + post-process), from a separate timing"""
    import torch
    st = pipe._static
    x = st["x"].clone()
    for _ in range(2):
        pipe.decode(x, "uint8")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    pipe.decode(x, "uint8")
    e1.record()
    torch.cuda.synchronize()
    t_vae = e0.elapsed_time(e1)
    e0.record()
    pipe.denoise(x * 0 + 1, st["ctx2"], pipe.scheduler.num_inference_steps)
    e1.record()
    torch.cuda.synchronize()
    t_den = e0.elapsed_time(e1)
    return t_vae / (t_vae + t_den)


# ---------------------------------------------------------------------------------------------------- patch-parallel arm
def run_patch_parallel(args):
    """BASELINE configs[4]: one request (batch `--batch`, default use 1) denoised by all ranks together, each on its band of
    latent rows (b200sd/patch.py).  A step of this arm = one whole image: `denoise_steps` UNet band steps + CFG/scheduler
    on the band, the all-gather of the final latent bands, and the VAE decode (on every rank, not split).  Strong scaling:
    the total work is fixed as N grows.  Rank 0 also runs the same request on its GPU alone and reports the PSNR between the
    two uint8 images."""
    import math

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29531")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    import b200sd
    from b200sd import ops
    from b200sd._randinit import random_state_dicts
    from b200sd.patch import BandComm, PatchParallelPipeline

    hw, B = args.size // 8, args.batch
    ucfg = b200sd.UNetConfig(sample_size=hw, prediction_type="v_prediction")
    vcfg = b200sd.VAEConfig()
    usd, vsd = random_state_dicts(ucfg, vcfg, 0)
    unet, vae = b200sd.B200UNet(usd, ucfg, dev), b200sd.B200VAEDecoder(vsd, vcfg, dev)
    del usd, vsd
    comm = BandComm()
    pipe = PatchParallelPipeline(unet, vae, b200sd.B200Scheduler(args.scheduler, ucfg.prediction_type, dev), comm)
    g = torch.Generator("cpu").manual_seed(1234)          # the same request on every rank
    lat = torch.randn(B, 4, hw, hw, generator=g)
    cond = torch.randn(B, 77, 1024, generator=g)
    unc = torch.randn(1, 77, 1024, generator=torch.Generator("cpu").manual_seed(7)).expand(B, -1, -1).contiguous()

    def one(time_steps=False):
        return pipe(prompt_embeds=cond, negative_prompt_embeds=unc, latents=lat, num_inference_steps=args.denoise_steps,
                    time_steps=time_steps).images

    for _ in range(max(args.warmup, 1)):
        one()
    torch.cuda.synchronize()
    # accounting pass, eager: a graph replay re-issues the captured launches and NCCL calls without passing through the
    # Python counters
    saved = (pipe.use_cuda_graph, pipe._graph)
    pipe.use_cuda_graph, pipe._graph = False, None
    n0, comm.bytes_gathered, comm.bytes_halo, comm.n_allreduce = ops.launch_count(), 0, 0, 0
    one()
    torch.cuda.synchronize()
    launches = ops.launch_count() - n0
    gathered, halo, n_ar = comm.bytes_gathered, comm.bytes_halo, comm.n_allreduce
    pipe.use_cuda_graph, pipe._graph = saved
    del saved   # (a second reference would keep the captured NCCL kernels alive past pipe.release())
    clocks = ClockSampler(local)
    dist.barrier()
    torch.cuda.synchronize()
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        img = one()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    dist.barrier()
    one(time_steps=True)                                   # per-step times (a synchronize per step: outside the timed region)
    step_ms = sorted(pipe.step_ms)[len(pipe.step_ms) // 2]
    t = torch.tensor([ms, step_ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, step_ms = float(t[0]), float(t[1])
    if rank == 0:
        single = b200sd.B200Pipeline(unet, vae, b200sd.B200Scheduler(args.scheduler, ucfg.prediction_type, dev))
        ref = single(prompt_embeds=cond, negative_prompt_embeds=unc, latents=lat, num_inference_steps=args.denoise_steps,
                     height=args.size, width=args.size, output_type="uint8").images
        e0.record()
        single.denoise(lat.to(dev).permute(0, 2, 3, 1).contiguous(), torch.cat([unc, cond]).to(dev).to(unet.K.op_dtype),
                       args.denoise_steps)
        e1.record()
        torch.cuda.synchronize()
        one_gpu_step_ms = e0.elapsed_time(e1) / args.denoise_steps
        mse = float(((img.float() - ref.float()) ** 2).mean())
        psnr = 99.0 if mse == 0 else 10.0 * math.log10(255.0 ** 2 / mse)
        value = B * args.steps / (ms * 1e-3)
        pk = peaks()
        tflop_img = args.denoise_steps * UNET_TFLOP.get(hw, 0) + VAE_TFLOP.get(hw, 0)
        line = {"metric": "sd21_images_per_sec", "value": round(value, 4), "unit": "images/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None,
                "dtype": {torch.float16: "fp16", torch.bfloat16: "bf16"}[unet.K.op_dtype], "data": "synthetic",
                "config": {"workload": f"SD2.1 {args.size}x{args.size}, {args.denoise_steps} {args.scheduler} steps, CFG 7.5, "
                                       f"{B} image(s) patch-parallel over {world} GPU(s) by latent rows (BASELINE configs[4])",
                           "size": args.size, "denoise_steps": args.denoise_steps, "batch": B,
                           "parallelism": f"pp{world} (row bands; NCCL all-gather of self-attention K|V into local memory, "
                                          f"send/recv halo rows, all-reduce of GroupNorm sums; VAE decode not split)"},
                "unet_step_ms": round(step_ms, 3), "unet_step_ms_one_gpu_graph": round(one_gpu_step_ms, 3),
                "per_step": {"allgather_bytes_received_per_gpu": gathered // args.denoise_steps,
                             "halo_bytes_received_per_gpu": halo // args.denoise_steps,
                             "groupnorm_allreduces": n_ar // args.denoise_steps},
                "psnr_vs_single_gpu_db": round(psnr, 2),
                "whole_image_tensor_roofline_frac": round(value * tflop_img / (world * pk["tflops"]), 4) if tflop_img else None,
                "gpu_launches": int(launches * args.steps), "clocks": clk}
        _emit(line)
    pipe.release()   # the captured step holds NCCL kernels: drop it before the group goes away
    dist.barrier()
    dist.destroy_process_group()


_REAL_STDOUT = None


def _emit(line: dict) -> None:
    """the ONE JSON line, on the process's real stdout"""
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.dup2(_REAL_STDOUT, 1)
    print(json.dumps(line), flush=True)


def main():
    global _REAL_STDOUT
    args = parse()
    # Libraries print to file descriptor 1 behind Python's back (NCCL's "NCCL version ..." banner at communicator
    # creation): everything but the JSON line goes to stderr, so stdout carries exactly one line.
    if not (args.profile_only or args.shape_profile):
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    elif args.patch_parallel:
        run_patch_parallel(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()

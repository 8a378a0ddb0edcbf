"""B200Scheduler -- drop-in for `pipe.scheduler` (the reference forces DDIMScheduler at app/run-sd.py:108;
BASELINE.json also names DPM-Solver++ and Euler).

All three are per-step LINEAR updates in (latent, guided model output, previous x0), so the host computes six
fp64 scalars per step, uploads the whole table once per `set_timesteps`, and one CUDA kernel
(`b200sd::cfg_sched_step`) does CFG + the update for any of them.  The diffusers surface
(`set_timesteps`, `timesteps`, `init_noise_sigma`, `scale_model_input`, `step(...)`, `order`, `config`) is kept.
"""
from __future__ import annotations

import math
from types import SimpleNamespace
from typing import List, Optional

import torch

from . import ops

KINDS = ("ddim", "dpm++2m", "euler")


def _alphas_cumprod() -> List[float]:
    n = 1000
    lo, hi = math.sqrt(0.00085), math.sqrt(0.012)
    out, acc = [], 1.0
    for i in range(n):
        beta = (lo + (hi - lo) * i / (n - 1)) ** 2
        acc *= 1.0 - beta
        out.append(acc)
    return out


def _norm_device(device) -> torch.device:
    """'cuda' -> 'cuda:<current>': torch.device('cuda') != torch.device('cuda:0'), and tensors always report an index"""
    d = torch.device(device)
    if d.type == "cuda" and d.index is None and torch.cuda.is_available():
        d = torch.device("cuda", torch.cuda.current_device())
    return d


class B200Scheduler:
    order = 1

    def __init__(self, kind: str = "ddim", prediction_type: str = "v_prediction", device="cuda"):
        if kind not in KINDS:
            raise ValueError(f"scheduler kind {kind!r} not in {KINDS}")
        if prediction_type not in ("v_prediction", "epsilon"):
            raise ValueError(prediction_type)
        self.kind, self.prediction_type = kind, prediction_type
        self.device = _norm_device(device)
        self.acp = _alphas_cumprod()
        self.config = SimpleNamespace(num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012,
                                      beta_schedule="scaled_linear", prediction_type=prediction_type,
                                      steps_offset=1, timestep_spacing="leading", clip_sample=False,
                                      set_alpha_to_one=False)
        self.init_noise_sigma = 1.0
        self.timesteps = torch.empty(0, dtype=torch.int64)
        self.num_inference_steps = 0
        self._rows: List[List[float]] = []
        self._in_scale: List[float] = []
        self._coef_dev: Optional[torch.Tensor] = None
        self.coef_generation = 0
        self._guidance = None
        self._step_of_t = {}
        self._x0_prev = None

    # ---- coefficient table ---------------------------------------------------------------------------------
    def _x0_coef(self, a: float):
        sa, sb = math.sqrt(a), math.sqrt(1.0 - a)
        return (sa, -sb) if self.prediction_type == "v_prediction" else (1.0 / sa, -sb / sa)

    def set_timesteps(self, num_inference_steps: int, device=None):
        n = int(num_inference_steps)
        if device is not None:
            self.device = _norm_device(device)
        ratio = 1000 // n
        ts = [i * ratio + 1 for i in range(n)][::-1]  # "leading" spacing, steps_offset = 1
        self.num_inference_steps = n
        self._ts = ts
        self.timesteps = torch.tensor(ts, dtype=torch.int64, device=self.device)
        self._step_of_t = {t: i for i, t in enumerate(ts)}
        acp = self.acp
        rows, in_scale = [], [1.0] * (n + 1)
        if self.kind == "ddim":
            self.init_noise_sigma = 1.0
            for t in ts:
                tp = t - ratio
                a, ap = acp[t], (acp[tp] if tp >= 0 else acp[0])
                sa, sb, sap, sbp = math.sqrt(a), math.sqrt(1 - a), math.sqrt(ap), math.sqrt(1 - ap)
                ax, am = self._x0_coef(a)
                if self.prediction_type == "v_prediction":
                    cx, cm = sap * sa + sbp * sb, sbp * sa - sap * sb
                else:
                    cx, cm = sap / sa, sbp - sap * sb / sa
                rows.append([cx, cm, 0.0, ax, am])
        else:
            sig = [math.sqrt((1 - acp[t]) / acp[t]) for t in ts] + [0.0]
            if self.kind == "euler":
                self.init_noise_sigma = math.sqrt(sig[0] ** 2 + 1.0)
                in_scale = [1.0 / math.sqrt(s * s + 1.0) for s in sig[:-1]] + [1.0]
                for i in range(n):
                    s, sn = sig[i], sig[i + 1]
                    if self.prediction_type == "v_prediction":
                        ax, am = 1.0 / (s * s + 1.0), -s / math.sqrt(s * s + 1.0)
                    else:
                        ax, am = 1.0, -s
                    k = (sn - s) / s
                    rows.append([1.0 + k * (1.0 - ax), -k * am, 0.0, ax, am])
            else:  # DPM-Solver++(2M), midpoint, lower_order_final, final sigma 0
                self.init_noise_sigma = 1.0
                alpha = lambda s: 1.0 / math.sqrt(s * s + 1.0)
                lam = lambda s: math.log(alpha(s)) - math.log(s * alpha(s))
                for i in range(n):
                    ax, am = self._x0_coef(acp[ts[i]])
                    if i == n - 1:
                        rows.append([ax, am, 0.0, ax, am])
                        continue
                    s_s, s_t = sig[i], sig[i + 1]
                    h = lam(s_t) - lam(s_s)
                    k0 = (s_t * alpha(s_t)) / (s_s * alpha(s_s))
                    k1 = -alpha(s_t) * math.expm1(-h)
                    if i == 0:
                        d0, dp = k1, 0.0
                    else:
                        r0 = (lam(s_s) - lam(sig[i - 1])) / h
                        d0, dp = k1 * (1 + 0.5 / r0), -k1 * 0.5 / r0
                    rows.append([k0 + d0 * ax, d0 * am, dp, ax, am])
        self._rows, self._in_scale = rows, in_scale
        self._guidance = None  # forces a re-upload; the device buffer itself is kept (CUDA graphs hold its address)
        self._x0_prev = None

    @property
    def needs_x0_prev(self) -> bool:
        return self.kind == "dpm++2m"

    def coef_table(self, guidance: float) -> torch.Tensor:
        """device fp32 [n_steps, 6] = {g, cx, cm, cp, ax, am}; uploaded once per (set_timesteps, guidance)."""
        if self._coef_dev is None or self._guidance != guidance:
            host = torch.tensor([[guidance] + r for r in self._rows], dtype=torch.float64).to(torch.float32)
            # The buffer is rewritten in place whenever the shape allows: captured CUDA graphs hold its address.
            if self._coef_dev is not None and self._coef_dev.shape == host.shape and self._coef_dev.device == self.device:
                self._coef_dev.copy_(host)
            else:
                self._coef_dev = host.to(self.device)
                self.coef_generation += 1  # a new address: graphs captured against the old buffer must be re-captured
            self._guidance = guidance
        return self._coef_dev

    def in_scale(self, i: int) -> float:
        return self._in_scale[i]

    # ---- diffusers surface -----------------------------------------------------------------------------------
    def scale_model_input(self, sample: torch.Tensor, timestep=None) -> torch.Tensor:
        if self.kind != "euler":
            return sample
        return sample * self._in_scale[self._index(timestep)]

    def _index(self, timestep) -> int:
        t = int(timestep.item() if torch.is_tensor(timestep) else timestep)
        return self._step_of_t[t]

    def step(self, model_output: torch.Tensor, timestep, sample: torch.Tensor, eta: float = 0.0,
             generator=None, return_dict: bool = True, **_unused):
        """Scheduler update for an already guided `model_output` (the stock diffusers pipeline does CFG itself):
        runs the fused kernel with g = 1 on (model_output, model_output)."""
        i = self._index(timestep)
        x = sample.to(device=self.device, dtype=torch.float32).contiguous().clone()
        m = model_output.to(device=self.device, dtype=torch.float32).contiguous()
        eps2 = torch.cat([m.reshape(-1), m.reshape(-1)])
        x0p = None
        if self.needs_x0_prev:
            if self._x0_prev is None or self._x0_prev.shape != x.shape:
                self._x0_prev = torch.zeros_like(x)
            x0p = self._x0_prev
        idx = torch.tensor([i], dtype=torch.int32, device=self.device)
        ops.cfg_sched_step(eps2, x, x0p, self.coef_table(1.0), idx)
        out = x.to(sample.dtype)
        return SimpleNamespace(prev_sample=out) if return_dict else (out,)

    def step_cfg(self, eps2: torch.Tensor, step_index: torch.Tensor, x: torch.Tensor,
                 x0_prev: Optional[torch.Tensor], guidance: float) -> None:
        """fast path used by B200Pipeline: eps2 = [uncond; cond] raw UNet output, x fp32 updated in place."""
        ops.cfg_sched_step(eps2, x, x0_prev, self.coef_table(guidance), step_index)

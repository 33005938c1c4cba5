"""Host side of the schedulers: integer timestep tables (bit-exact) and the per-step coefficient table consumed on
the device by `b200sd_cfg_scheduler_step` / `b200sd_prep_unet_input` / `b200sd_timestep_embedding`.

Replaces `scheduler.set_timesteps`, `scheduler.scale_model_input` and the scalar half of `scheduler.step` for
DDIMScheduler, PNDMScheduler (skip_prk_steps, the SD1.5 repo default) and EulerDiscreteScheduler, as driven by
`pipe(..., num_inference_steps=N)` (reference configmap.yaml:107).  All schedules:
scaled_linear betas, epsilon prediction, "leading" spacing, set_alpha_to_one=False.  Scalar coefficients are
computed in fp32 with torch, in the same operation order as upstream, then uploaded once per call.
"""
from __future__ import annotations

import numpy as np
import torch

from . import lib as L

KIND = {"ddim": 0, "euler": 1, "pndm": 2}


def alphas_cumprod(beta_start=0.00085, beta_end=0.012, n=1000) -> torch.Tensor:
    betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, n, dtype=torch.float32) ** 2
    return torch.cumprod(1.0 - betas, dim=0)


class StepSchedule:
    """timesteps: int64 list fed to the UNet (length = number of UNet evaluations); table: f32 [evals, 16];
    init_noise_sigma: factor on the initial N(0,1) latents."""

    def __init__(self, kind, timesteps, table, init_noise_sigma):
        self.kind, self.timesteps, self.table, self.init_noise_sigma = kind, timesteps, table, init_noise_sigma

    @property
    def num_evals(self):
        return len(self.timesteps)


def make_schedule(kind: str, num_inference_steps: int, *, steps_offset=1, beta_start=0.00085, beta_end=0.012,
                  num_train_timesteps=1000, **_) -> StepSchedule:
    kind = kind.lower()
    n, T = num_inference_steps, num_train_timesteps
    if n < 1 or n > T:
        raise ValueError(f"num_inference_steps must be in [1, {T}] (got {n})")
    ac = alphas_cumprod(beta_start, beta_end, T)
    final_ac = ac[0]
    ratio = T // n
    W = L.ST_WIDTH

    def a(t):
        return ac[t] if t >= 0 else final_ac

    if kind == "ddim":
        ts = ((np.arange(0, n) * ratio).round()[::-1].copy().astype(np.int64) + steps_offset).tolist()
        table = torch.zeros(n, W, dtype=torch.float32)
        for i, t in enumerate(ts):
            a_t, a_p = a(t), a(t - ratio)
            table[i, L.ST_TIMESTEP] = float(t)
            table[i, L.ST_IN_SCALE] = 1.0
            table[i, L.ST_KIND] = 0
            table[i, L.ST_C0] = (1 - a_t) ** 0.5
            table[i, L.ST_C1] = a_t ** 0.5
            table[i, L.ST_C2] = a_p ** 0.5
            table[i, L.ST_C3] = (1 - a_p) ** 0.5
        return StepSchedule(kind, ts, table, 1.0)

    if kind == "pndm":
        _t = (np.arange(0, n) * ratio).round() + steps_offset
        ts = np.concatenate([_t[:-1], _t[-2:-1], _t[-1:]])[::-1].copy().astype(np.int64).tolist()
        table = torch.zeros(len(ts), W, dtype=torch.float32)
        n_ets = 0
        for counter, t in enumerate(ts):
            prev = t - ratio
            push = counter != 1
            if push:
                n_ets = min(n_ets + 1, 4)
            else:
                prev, t_eff = t, t + ratio
            t_used = t if push else t_eff
            mode = 0
            w = [0.0, 0.0, 0.0, 0.0]
            if n_ets == 1 and counter == 0:
                mode = 1
            elif n_ets == 1 and counter == 1:
                mode = 2
            elif n_ets == 2:
                w = [3 / 2, -1 / 2, 0.0, 0.0]
            elif n_ets == 3:
                w = [23 / 12, -16 / 12, 5 / 12, 0.0]
            else:
                w = [55 / 24, -59 / 24, 37 / 24, -9 / 24]
            a_t, a_p = a(t_used), a(prev)
            b_t, b_p = 1 - a_t, 1 - a_p
            sample_coeff = (a_p / a_t) ** 0.5
            denom = a_t * b_p ** 0.5 + (a_t * b_t * a_p) ** 0.5
            table[counter, L.ST_TIMESTEP] = float(t)
            table[counter, L.ST_IN_SCALE] = 1.0
            table[counter, L.ST_KIND] = 2
            table[counter, L.ST_C0] = sample_coeff
            table[counter, L.ST_C1] = (a_p - a_t) / denom
            for j in range(4):
                table[counter, L.ST_W0 + j] = w[j]
            table[counter, L.ST_PNDM_MODE] = mode
            table[counter, L.ST_PUSH] = 1.0 if push else 0.0
        return StepSchedule(kind, ts, table, 1.0)

    if kind == "euler":
        tsf = (np.arange(0, n) * ratio).round()[::-1].copy().astype(np.float32) + steps_offset
        sig = (((1 - ac) / ac) ** 0.5).numpy()
        sig = np.interp(tsf, np.arange(0, len(sig)), sig)
        sig = torch.from_numpy(np.concatenate([sig, [0.0]]).astype(np.float32))
        table = torch.zeros(n, W, dtype=torch.float32)
        for i in range(n):
            s = sig[i]
            table[i, L.ST_TIMESTEP] = float(tsf[i])
            table[i, L.ST_IN_SCALE] = 1.0 / ((s ** 2 + 1) ** 0.5)
            table[i, L.ST_KIND] = 1
            table[i, L.ST_C0] = sig[i + 1] - s
            table[i, L.ST_C1] = s
        init = float((sig.max() ** 2 + 1) ** 0.5)
        return StepSchedule(kind, [int(t) for t in tsf], table, init)

    raise ValueError(f"unknown scheduler {kind!r} (ddim | pndm | euler)")

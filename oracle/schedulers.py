"""Oracle schedulers (fp32 torch / numpy) — restate diffusers==0.30.2 `schedulers/scheduling_ddim.py`,
`scheduling_pndm.py`, `scheduling_euler_discrete.py` (pinned at reference configmap.yaml:13).  The reference never
names a scheduler (configmap.yaml:41): the model repo's default applies — PNDM with skip_prk_steps, steps_offset 1
for runwayml/stable-diffusion-v1-5; BASELINE.json measures with DDIM; SDXL ships Euler.
All three: scaled_linear betas, epsilon prediction, "leading" timestep spacing, set_alpha_to_one=False.
Timestep tables are integer-exact; known answers in SURVEY.md App. C are asserted in tests/test_oracle.py.
TEST INFRASTRUCTURE — see oracle/__init__.py."""
from __future__ import annotations

import numpy as np
import torch


def alphas_cumprod(beta_start=0.00085, beta_end=0.012, n=1000) -> torch.Tensor:
    betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, n, dtype=torch.float32) ** 2
    return torch.cumprod(1.0 - betas, dim=0)


class DDIMScheduler:
    order = 1
    init_noise_sigma = 1.0

    def __init__(self, steps_offset=1, beta_start=0.00085, beta_end=0.012, num_train_timesteps=1000, **_):
        self.n_train = num_train_timesteps
        self.steps_offset = steps_offset
        self.alphas_cumprod = alphas_cumprod(beta_start, beta_end, num_train_timesteps)
        self.final_alpha_cumprod = self.alphas_cumprod[0]  # set_alpha_to_one=False

    def set_timesteps(self, n):
        self.num_inference_steps = n
        ratio = self.n_train // n
        ts = (np.arange(0, n) * ratio).round()[::-1].copy().astype(np.int64) + self.steps_offset
        self.timesteps = torch.from_numpy(ts)

    def scale_model_input(self, x, t):
        return x

    def step(self, eps, t, x):
        t = int(t)
        prev = t - self.n_train // self.num_inference_steps
        a_t = self.alphas_cumprod[t]
        a_p = self.alphas_cumprod[prev] if prev >= 0 else self.final_alpha_cumprod
        x0 = (x - (1 - a_t) ** 0.5 * eps) / a_t ** 0.5
        direction = (1 - a_p) ** 0.5 * eps  # eta = 0 -> std_dev_t = 0
        return a_p ** 0.5 * x0 + direction


class PNDMScheduler:
    """PLMS branch only (skip_prk_steps=True)."""
    order = 1
    init_noise_sigma = 1.0

    def __init__(self, steps_offset=1, beta_start=0.00085, beta_end=0.012, num_train_timesteps=1000, **_):
        self.n_train = num_train_timesteps
        self.steps_offset = steps_offset
        self.alphas_cumprod = alphas_cumprod(beta_start, beta_end, num_train_timesteps)
        self.final_alpha_cumprod = self.alphas_cumprod[0]

    def set_timesteps(self, n):
        self.num_inference_steps = n
        ratio = self.n_train // n
        _t = (np.arange(0, n) * ratio).round() + self.steps_offset
        plms = np.concatenate([_t[:-1], _t[-2:-1], _t[-1:]])[::-1].copy()
        self.timesteps = torch.from_numpy(plms.astype(np.int64))
        self.ets = []
        self.counter = 0
        self.cur_sample = None

    def scale_model_input(self, x, t):
        return x

    def step(self, eps, t, x):
        t = int(t)
        prev = t - self.n_train // self.num_inference_steps
        if self.counter != 1:
            self.ets = self.ets[-3:]
            self.ets.append(eps)
        else:
            prev = t
            t = t + self.n_train // self.num_inference_steps
        if len(self.ets) == 1 and self.counter == 0:
            e = eps
            self.cur_sample = x
        elif len(self.ets) == 1 and self.counter == 1:
            e = (eps + self.ets[-1]) / 2
            x = self.cur_sample
            self.cur_sample = None
        elif len(self.ets) == 2:
            e = (3 * self.ets[-1] - self.ets[-2]) / 2
        elif len(self.ets) == 3:
            e = (23 * self.ets[-1] - 16 * self.ets[-2] + 5 * self.ets[-3]) / 12
        else:
            e = (1 / 24) * (55 * self.ets[-1] - 59 * self.ets[-2] + 37 * self.ets[-3] - 9 * self.ets[-4])
        out = self._prev_sample(x, t, prev, e)
        self.counter += 1
        return out

    def _prev_sample(self, x, t, prev, e):
        a_t = self.alphas_cumprod[t]
        a_p = self.alphas_cumprod[prev] if prev >= 0 else self.final_alpha_cumprod
        b_t, b_p = 1 - a_t, 1 - a_p
        sample_coeff = (a_p / a_t) ** 0.5
        denom = a_t * b_p ** 0.5 + (a_t * b_t * a_p) ** 0.5
        return sample_coeff * x - (a_p - a_t) * e / denom


class EulerDiscreteScheduler:
    order = 1

    def __init__(self, steps_offset=1, beta_start=0.00085, beta_end=0.012, num_train_timesteps=1000, **_):
        self.n_train = num_train_timesteps
        self.steps_offset = steps_offset
        self.alphas_cumprod = alphas_cumprod(beta_start, beta_end, num_train_timesteps)
        sig = (((1 - self.alphas_cumprod) / self.alphas_cumprod) ** 0.5).flip(0)
        self.sigmas = torch.cat([sig, torch.zeros(1)])
        self.init_noise_sigma = float((self.sigmas.max() ** 2 + 1) ** 0.5)

    def set_timesteps(self, n):
        self.num_inference_steps = n
        ratio = self.n_train // n
        ts = (np.arange(0, n) * ratio).round()[::-1].copy().astype(np.float32) + self.steps_offset
        sig = (((1 - self.alphas_cumprod) / self.alphas_cumprod) ** 0.5).numpy()
        sig = np.interp(ts, np.arange(0, len(sig)), sig)
        sig = np.concatenate([sig, [0.0]]).astype(np.float32)
        self.sigmas = torch.from_numpy(sig)
        self.timesteps = torch.from_numpy(ts)
        self.init_noise_sigma = float((self.sigmas.max() ** 2 + 1) ** 0.5)  # "leading" spacing
        self._idx = 0

    def scale_model_input(self, x, t):
        s = self.sigmas[self._idx]
        return x / ((s ** 2 + 1) ** 0.5)

    def step(self, eps, t, x):
        x = x.to(torch.float32)
        s = self.sigmas[self._idx]
        x0 = x - s * eps
        derivative = (x - x0) / s
        dt = self.sigmas[self._idx + 1] - s
        self._idx += 1
        return x + derivative * dt


def make_scheduler(cfg: dict, kind: str | None = None):
    kind = (kind or cfg["type"]).lower()
    args = {k: v for k, v in cfg.items() if k != "type"}
    return {"ddim": DDIMScheduler, "pndm": PNDMScheduler, "euler": EulerDiscreteScheduler}[kind](**args)

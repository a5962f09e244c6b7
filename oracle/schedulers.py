"""Oracle: restated diffusers DDIMScheduler / EulerDiscreteScheduler as the reference configures them.

TEST INFRASTRUCTURE ONLY.  Follows SURVEY.md Appendix B.2 / B.3; the configuration comes from
/root/reference/experiments/05_grid_search_ablation.py:136-141 (DDIM: beta_schedule="linear",
steps_offset=1, clip_sample=False on SD1.5's scheduler config) and
/root/reference/experiments/01_baseline_generation.py:76-80 (Euler: timestep_spacing="linspace",
beta_schedule="linear").  Formulas: /root/reference/docs/01_diffusion_fundamentals.md:109-136.
Written op-by-op the way diffusers does it (including the per-step table lookup), no fusion.
"""
from __future__ import annotations

import numpy as np
import torch


def _alphas_cumprod(num_train: int = 1000, beta_start: float = 0.00085, beta_end: float = 0.012,
                    schedule: str = "linear") -> torch.Tensor:
    if schedule == "linear":
        betas = torch.linspace(beta_start, beta_end, num_train, dtype=torch.float32)
    elif schedule == "scaled_linear":
        betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_train, dtype=torch.float32) ** 2
    else:
        raise ValueError(schedule)
    return torch.cumprod(1.0 - betas, dim=0)


class OracleDDIM:
    """timestep_spacing='leading', set_alpha_to_one=False, prediction_type='epsilon', eta=0."""
    init_noise_sigma = 1.0
    order = 1

    def __init__(self, beta_schedule: str = "linear", steps_offset: int = 1, num_train: int = 1000):
        self.num_train = num_train
        self.steps_offset = steps_offset
        self.alphas_cumprod = _alphas_cumprod(num_train, schedule=beta_schedule)
        self.final_alpha_cumprod = self.alphas_cumprod[0]
        self.timesteps = None

    def set_timesteps(self, n: int):
        self.num_inference_steps = n
        ratio = self.num_train // n
        ts = (np.arange(0, n) * ratio).round()[::-1].copy().astype(np.int64) + self.steps_offset
        self.timesteps = torch.from_numpy(ts)

    def scale_model_input(self, x, t):
        return x

    def step(self, eps: torch.Tensor, t: int, x: torch.Tensor) -> torch.Tensor:
        t = int(t)
        prev = t - self.num_train // self.num_inference_steps
        a_t = self.alphas_cumprod[t]
        a_p = self.alphas_cumprod[prev] if prev >= 0 else self.final_alpha_cumprod
        beta_t = 1 - a_t
        x0 = (x - beta_t ** 0.5 * eps) / a_t ** 0.5
        direction = (1 - a_p) ** 0.5 * eps
        return a_p ** 0.5 * x0 + direction


class OracleEuler:
    """epsilon prediction, no churn (s_churn=0), interpolation 'linear'.  timestep_spacing='linspace' is the
    reference's configuration (01_baseline_generation.py:76-80); 'leading' (+steps_offset) and 'trailing' are what
    EulerDiscreteScheduler.from_config(pipe.scheduler.config) yields without the override."""
    order = 1

    def __init__(self, beta_schedule: str = "linear", num_train: int = 1000, timestep_spacing: str = "linspace",
                 steps_offset: int = 0):
        self.num_train = num_train
        self.timestep_spacing, self.steps_offset = timestep_spacing, steps_offset
        ac = _alphas_cumprod(num_train, schedule=beta_schedule)
        self.base_sigmas = (((1 - ac) / ac) ** 0.5).numpy().astype(np.float64)
        self.sigmas = torch.from_numpy(np.concatenate([self.base_sigmas[::-1], [0.0]]).astype(np.float32))
        self.timesteps = None

    def set_timesteps(self, n: int):
        if self.timestep_spacing == "linspace":
            ts = np.linspace(0, self.num_train - 1, n, dtype=np.float32)[::-1].copy()
        elif self.timestep_spacing == "leading":
            ratio = self.num_train // n
            ts = (np.arange(0, n) * ratio).round()[::-1].copy().astype(np.float32) + self.steps_offset
        elif self.timestep_spacing == "trailing":
            ts = (np.arange(self.num_train, 0, -self.num_train / n)).round().astype(np.float32) - 1
        else:
            raise ValueError(self.timestep_spacing)
        sig = np.interp(ts, np.arange(0, len(self.base_sigmas)), self.base_sigmas)
        sig = np.concatenate([sig, [0.0]]).astype(np.float32)
        self.sigmas = torch.from_numpy(sig)
        self.timesteps = torch.from_numpy(ts)
        self.step_index = 0

    @property
    def init_noise_sigma(self):
        m = float(self.sigmas.max())
        if self.timestep_spacing in ("linspace", "trailing"):
            return m
        return (m ** 2 + 1) ** 0.5

    def scale_model_input(self, x, t):
        s = self.sigmas[self.step_index]
        return x / ((s ** 2 + 1) ** 0.5)

    def step(self, eps, t, x):
        s = self.sigmas[self.step_index]
        x32 = x.to(torch.float32)
        pred_x0 = x32 - s * eps.to(torch.float32)
        derivative = (x32 - pred_x0) / s
        dt = self.sigmas[self.step_index + 1] - s
        self.step_index += 1
        return (x32 + derivative * dt).to(x.dtype)


def cfg_combine(noise_pred: torch.Tensor, guidance_scale: float) -> torch.Tensor:
    """u + g*(c-u), uncond row first (/root/reference/docs/01_diffusion_fundamentals.md:182-186)."""
    u, c = noise_pred.chunk(2)
    return u + guidance_scale * (c - u)


@torch.no_grad()
def denoise_loop(unet, scheduler, latents, text_embeds, guidance_scale, num_steps, on_step=None):
    """The loop body of AnimateDiffPipeline.__call__.  latents [1,4,F,h,w];
    text_embeds [2,77,768] with the negative (uncond) row first.  Returns final latents."""
    scheduler.set_timesteps(num_steps)
    latents = latents * scheduler.init_noise_sigma
    for i, t in enumerate(scheduler.timesteps):
        x_in = torch.cat([latents] * 2)
        x_in = scheduler.scale_model_input(x_in, t)
        eps = unet(x_in, t, encoder_hidden_states=text_embeds).sample
        guided = cfg_combine(eps, guidance_scale)
        latents = scheduler.step(guided, t, latents)
        if on_step is not None:
            on_step(i, t, eps, latents)
    return latents

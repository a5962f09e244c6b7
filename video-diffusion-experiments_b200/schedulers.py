"""DDIMScheduler / EulerDiscreteScheduler with the diffusers surface the reference uses
(/root/reference/experiments/05_grid_search_ablation.py:136-141, 01_baseline_generation.py:76-80):
``from_config(cfg, **overrides)``, ``.config``, ``set_timesteps``, ``timesteps``, ``init_noise_sigma``,
``scale_model_input``, ``step(...).prev_sample``, ``order``.

B200-first difference: ``set_timesteps`` precomputes, in fp64 on the host, one (c_x, c_eps) pair per step so that
the whole update is  x' = c_x*x + c_eps*eps  -- no per-step table lookup with a device scalar, no host sync --
and the pipeline fuses it with the classifier-free-guidance combine in one kernel (b200ad_cfg_step).
Timestep indexing is integer-exact with diffusers.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, Optional

import numpy as np
import torch


class FrozenConfig(dict):
    """Mapping with attribute access (stands in for diffusers' FrozenDict)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e


@dataclass
class SchedulerOutput:
    prev_sample: torch.Tensor


def _betas(cfg) -> torch.Tensor:
    n, b0, b1 = cfg["num_train_timesteps"], cfg["beta_start"], cfg["beta_end"]
    if cfg["beta_schedule"] == "linear":
        return torch.linspace(b0, b1, n, dtype=torch.float32)
    if cfg["beta_schedule"] == "scaled_linear":
        return torch.linspace(b0 ** 0.5, b1 ** 0.5, n, dtype=torch.float32) ** 2
    raise NotImplementedError(f"beta_schedule={cfg['beta_schedule']!r}")


class _SchedulerBase:
    order = 1
    _defaults: Dict[str, Any] = {}

    def __init__(self, **kwargs):
        cfg = dict(self._defaults)
        unknown = {k: v for k, v in kwargs.items() if k not in cfg}
        cfg.update({k: v for k, v in kwargs.items() if k in cfg})
        cfg["_class_name"] = type(self).__name__
        cfg["_ignored"] = sorted(unknown)          # keys of a foreign scheduler config (e.g. PNDM's skip_prk_steps)
        self.config = FrozenConfig(cfg)
        self.timesteps: Optional[torch.Tensor] = None
        self.num_inference_steps: Optional[int] = None
        self._setup()

    @classmethod
    def from_config(cls, config, **overrides):
        cfg = {k: v for k, v in dict(config).items() if not k.startswith("_")}
        cfg.update(overrides)
        return cls(**cfg)

    # fused-path interface -------------------------------------------------------------------------
    def step_coefficients(self, i: int):
        """(c_x, c_eps, next_in_scale) of step i: x' = c_x*x + c_eps*eps; next model input = x'*next_in_scale."""
        return self._cx[i], self._ce[i], self._next_scale[i]

    def first_in_scale(self) -> float:
        """scale_model_input factor of the first step (1 for DDIM)."""
        return 1.0

    def _step_index_of(self, timestep) -> int:
        t = float(timestep)
        ts = self._timesteps_host
        idx = np.nonzero(np.isclose(ts, t))[0]
        if len(idx) == 0:
            raise ValueError(f"timestep {t} is not in the current schedule")
        return int(idx[0])

    def step(self, model_output, timestep, sample, eta: float = 0.0, use_clipped_model_output=False, generator=None,
             variance_noise=None, return_dict: bool = True, **kw):
        if self.num_inference_steps is None:
            raise ValueError("call set_timesteps() before step()")
        if eta != 0.0:
            raise NotImplementedError("only the deterministic eta=0 update is implemented")
        i = self._step_index_of(timestep)
        cx, ce, _ = self.step_coefficients(i)
        from .ops import get_ops
        o = get_ops("fp32")
        x = sample.float().contiguous()
        e = model_output.float().contiguous()
        out = o.cfg_step(e, e, x, 0.0, cx, ce).to(sample.dtype)      # g = 0: guided == model_output
        self._advance(i)
        return SchedulerOutput(prev_sample=out) if return_dict else (out,)

    def _advance(self, i):
        pass


class DDIMScheduler(_SchedulerBase):
    """epsilon-prediction DDIM, eta = 0 (formulas: /root/reference/docs/01_diffusion_fundamentals.md:109-124)."""
    _defaults = dict(num_train_timesteps=1000, beta_start=0.0001, beta_end=0.02, beta_schedule="linear",
                     trained_betas=None, clip_sample=True, set_alpha_to_one=True, steps_offset=0,
                     prediction_type="epsilon", thresholding=False, timestep_spacing="leading",
                     rescale_betas_zero_snr=False)

    def _setup(self):
        c = self.config
        if c.prediction_type != "epsilon" or c.thresholding or c.rescale_betas_zero_snr:
            raise NotImplementedError("DDIMScheduler: only epsilon prediction without thresholding is implemented")
        if c.clip_sample:
            raise NotImplementedError("DDIMScheduler: clip_sample=True is not on the reference's path "
                                      "(05_grid_search_ablation.py:140 sets clip_sample=False)")
        self.betas = _betas(c)
        self.alphas_cumprod = torch.cumprod(1.0 - self.betas, dim=0)
        self.final_alpha_cumprod = torch.tensor(1.0) if c.set_alpha_to_one else self.alphas_cumprod[0]
        self.init_noise_sigma = 1.0

    def set_timesteps(self, num_inference_steps: int, device=None):
        c = self.config
        n, T = int(num_inference_steps), c.num_train_timesteps
        if n > T:
            raise ValueError("num_inference_steps > num_train_timesteps")
        self.num_inference_steps = n
        if c.timestep_spacing == "leading":
            ratio = T // n
            ts = (np.arange(0, n) * ratio).round()[::-1].copy().astype(np.int64) + c.steps_offset
        elif c.timestep_spacing == "trailing":
            ts = np.round(np.arange(T, 0, -T / n)).astype(np.int64) - 1
        elif c.timestep_spacing == "linspace":
            ts = np.linspace(0, T - 1, n).round()[::-1].copy().astype(np.int64)
        else:
            raise ValueError(c.timestep_spacing)
        self._timesteps_host = ts
        self.timesteps = torch.from_numpy(ts).to(device) if device is not None else torch.from_numpy(ts)
        ac = self.alphas_cumprod.double().numpy()
        final = float(self.final_alpha_cumprod)
        self._cx, self._ce, self._next_scale, self._prev_t = [], [], [], []
        for t in ts:
            p = int(t) - T // n
            a_t = ac[int(t)]
            a_p = ac[p] if p >= 0 else final
            self._prev_t.append(p)
            self._cx.append(math.sqrt(a_p / a_t))
            self._ce.append(math.sqrt(1.0 - a_p) - math.sqrt(a_p * (1.0 - a_t) / a_t))
            self._next_scale.append(1.0)

    def scale_model_input(self, sample, timestep=None):
        return sample


class EulerDiscreteScheduler(_SchedulerBase):
    """epsilon-prediction Euler, no churn (formulas: /root/reference/docs/01_diffusion_fundamentals.md:130-136)."""
    _defaults = dict(num_train_timesteps=1000, beta_start=0.0001, beta_end=0.02, beta_schedule="linear",
                     trained_betas=None, prediction_type="epsilon", interpolation_type="linear",
                     use_karras_sigmas=False, timestep_spacing="linspace", steps_offset=0,
                     rescale_betas_zero_snr=False)

    def _setup(self):
        c = self.config
        if c.prediction_type != "epsilon" or c.use_karras_sigmas or c.interpolation_type != "linear":
            raise NotImplementedError("EulerDiscreteScheduler: only epsilon / linear interpolation is implemented")
        self.betas = _betas(c)
        self.alphas_cumprod = torch.cumprod(1.0 - self.betas, dim=0)
        # diffusers evaluates this in fp32 on the torch table, then interpolates in fp64 (np.interp)
        ac = self.alphas_cumprod
        self._base_sigmas = (((1 - ac) / ac) ** 0.5).numpy().astype(np.float64)
        self.sigmas = torch.from_numpy(np.concatenate([self._base_sigmas[::-1], [0.0]]).astype(np.float32))
        self._step_index = None

    @property
    def init_noise_sigma(self):
        s = float(self.sigmas.max())
        if self.config.timestep_spacing in ("linspace", "trailing"):
            return s
        return (s ** 2 + 1) ** 0.5

    def set_timesteps(self, num_inference_steps: int, device=None):
        c = self.config
        n, T = int(num_inference_steps), c.num_train_timesteps
        self.num_inference_steps = n
        if c.timestep_spacing == "linspace":
            ts = np.linspace(0, T - 1, n, dtype=np.float32)[::-1].copy()
        elif c.timestep_spacing == "leading":
            ts = ((np.arange(0, n) * (T // n)).round()[::-1].copy().astype(np.float32)) + c.steps_offset
        elif c.timestep_spacing == "trailing":
            ts = (np.arange(T, 0, -T / n)).round().astype(np.float32) - 1
        else:
            raise ValueError(c.timestep_spacing)
        sig = np.interp(ts, np.arange(0, T), self._base_sigmas)
        sig = np.concatenate([sig, [0.0]]).astype(np.float32)
        self.sigmas = torch.from_numpy(sig)
        self._timesteps_host = ts
        self.timesteps = torch.from_numpy(ts).to(device) if device is not None else torch.from_numpy(ts)
        s64 = sig.astype(np.float64)
        self._cx = [1.0] * n
        self._ce = [float(s64[i + 1] - s64[i]) for i in range(n)]
        self._next_scale = [float(1.0 / math.sqrt(s64[i + 1] ** 2 + 1.0)) for i in range(n)]
        self._step_index = 0

    def first_in_scale(self) -> float:
        return float(1.0 / math.sqrt(float(self.sigmas[0]) ** 2 + 1.0))

    def scale_model_input(self, sample, timestep=None):
        i = self._step_index if timestep is None else self._step_index_of(timestep)
        s = float(self.sigmas[i])
        return sample / ((s ** 2 + 1) ** 0.5)

    def _advance(self, i):
        self._step_index = i + 1


# SD1.5 ships a PNDM scheduler config; the reference rebuilds DDIM / Euler from it (05:136-141, 01:76-80).
SD15_SCHEDULER_CONFIG = FrozenConfig(
    _class_name="PNDMScheduler", num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012,
    beta_schedule="scaled_linear", trained_betas=None, skip_prk_steps=True, set_alpha_to_one=False,
    prediction_type="epsilon", timestep_spacing="leading", steps_offset=1, clip_sample=False)


class PNDMPlaceholder:
    """What ``from_pretrained`` leaves in ``pipe.scheduler``: SD1.5 ships a PNDM config, and every reference
    script replaces it via ``XScheduler.from_config(pipe.scheduler.config, ...)`` before generating
    (05_grid_search_ablation.py:136-141, 01_baseline_generation.py:76-80).  Only ``.config`` is usable."""

    def __init__(self, config):
        self.config = config

    def __getattr__(self, k):
        raise NotImplementedError(
            f"PNDMScheduler.{k}: PNDM is not implemented; assign DDIMScheduler.from_config(pipe.scheduler.config, ...)"
            " or EulerDiscreteScheduler.from_config(...) as the reference scripts do")

"""AnimateDiffPipeline / MotionAdapter with the diffusers surface the reference drives
(/root/reference/experiments/05_grid_search_ablation.py:121-169, 01_baseline_generation.py:60-108):
``from_pretrained``, ``.scheduler`` swap, ``enable_vae_slicing`` / ``enable_model_cpu_offload`` / ``to``, and
``pipe(prompt=, negative_prompt=, num_frames=, guidance_scale=, num_inference_steps=, height=, width=,
generator=).frames[0]`` -> list of PIL images.

B200-first loop: one denoise step (both CFG halves of the UNet + guidance combine + scheduler update + next
step's model input) is captured ONCE per (F, h, w) shape as a CUDA graph and replayed per step after a 32-byte
parameter upload (timestep and the scheduler's per-step coefficients); there is no per-step host sync, no
``torch.cat``, no table lookup with a device scalar.  With ``cfg_group`` (2 ranks) each rank runs one CFG half
and the halves meet in one all-gather of the noise predictions per step.
"""
from __future__ import annotations

import hashlib
import os
import sys
from dataclasses import dataclass
from typing import List, Optional, Union

import numpy as np
import torch

from . import schedulers as _sched
from .ops import get_ops
from .unet import UNetMotionConfig, UNetMotionModel
from .vae import AutoencoderKLDecoder


def _log(msg: str):
    print(f"[b200ad] {msg}", file=sys.stderr, flush=True)      # stdout stays clean (bench.py prints one JSON line)


@dataclass
class AnimateDiffPipelineOutput:
    frames: Union[List[List["PIL.Image.Image"]], np.ndarray, torch.Tensor]


class MotionAdapter:
    """Handle for the motion-module weights.  ``from_pretrained`` resolves a local diffusers-layout directory
    (``diffusion_pytorch_model.safetensors``); without one (no network on this box) the pipeline keeps the
    seeded random initialisation of its motion modules and says so."""

    def __init__(self, state_dict=None, source: str = ""):
        self.state_dict_ = state_dict
        self.source = source

    @classmethod
    def from_pretrained(cls, name_or_path: str, torch_dtype=None, **kw):
        sd = _load_local_state_dict(name_or_path, "")
        if sd is None:
            _log(f"MotionAdapter '{name_or_path}': no local checkpoint -> motion modules keep seeded random init")
        return cls(sd, name_or_path)

    def to(self, *a, **kw):
        return self


def _load_local_state_dict(path: str, sub: str):
    d = os.path.join(path, sub) if sub else path
    f = os.path.join(d, "diffusion_pytorch_model.safetensors")
    if not os.path.isfile(f):
        return None
    from safetensors.torch import load_file
    return load_file(f)


def synthetic_text_embeds(prompt: str, length: int = 77, dim: int = 768) -> torch.Tensor:
    """Deterministic stand-in for CLIPTextModel(prompt) when no CLIP weights exist on the box: a unit-variance
    [77, 768] tensor seeded by the prompt string (the text encoder is off the denoise path)."""
    seed = int.from_bytes(hashlib.sha256(prompt.encode("utf-8")).digest()[:8], "little") & ((1 << 63) - 1)
    g = torch.Generator().manual_seed(seed)
    return torch.randn(length, dim, generator=g)


class DenoisePlan:
    """Static buffers + (optionally) the captured CUDA graph of one denoise step for a fixed (B, F, h, w)."""

    def __init__(self, unet: UNetMotionModel, B: int, F: int, h: int, w: int, use_graph: bool, cfg_group=None):
        self.unet, self.B, self.F, self.h, self.w = unet, B, F, h, w
        self.use_graph = use_graph
        self.group = cfg_group
        self.split = cfg_group is not None
        dev = "cuda"
        nb = 1 if self.split else B                       # UNet batch on this rank
        self.model_in = torch.zeros(nb, 4, F, h, w, device=dev)
        self.latents = torch.zeros(1, 4, F, h, w, device=dev)
        self.text = torch.zeros(nb, 77, unet.cfg.cross_attention_dim, device=dev)
        self.params = torch.zeros(8, device=dev)          # [t, t, -, -, g, cx, ce, in_scale]
        self.eps_all = torch.zeros(2, 4, F, h, w, device=dev) if self.split else None
        self.table, self._table_key = None, None
        self.graph = None
        self.kv = None
        self.o32 = get_ops("fp32")

    # one denoise step on the current stream
    def _body(self):
        nb = self.model_in.shape[0]
        eps = self.unet(self.model_in, self.params[:nb], encoder_hidden_states=self.text, text_kv=self.kv).sample
        if self.split:
            torch.distributed.all_gather_into_tensor(self.eps_all, eps, group=self.group)
            eu, ec = self.eps_all[0], self.eps_all[1]
        else:
            eu, ec = eps[0], eps[self.B - 1]
        self.o32.cfg_step_dev(eu, ec, self.latents, self.params[4:8], self.latents, self.model_in)

    def load(self, latents, text, scheduler, guidance_scale: float):
        """Start a video: initial latents [1,4,F,h,w] (already * init_noise_sigma), text rows, schedule table."""
        n = scheduler.num_inference_steps
        key = (id(scheduler), type(scheduler).__name__, n, float(guidance_scale), id(scheduler._timesteps_host))
        if key != self._table_key:
            tab = torch.zeros(n, 8)
            ts = scheduler._timesteps_host
            for i in range(n):
                cx, ce, nxt = scheduler.step_coefficients(i)
                tab[i, 0] = tab[i, 1] = float(ts[i])
                tab[i, 4], tab[i, 5], tab[i, 6], tab[i, 7] = (guidance_scale if self.B == 2 else 0.0), cx, ce, nxt
            self.table, self._table_key = tab.pin_memory(), key
        self.latents.copy_(latents.reshape(self.latents.shape).float(), non_blocking=True)
        self.text.copy_(text.float(), non_blocking=True)
        self.kv = self.unet.text_kv(self.text, out=self.kv)        # in place: the captured graph reads these
        # first model input = latents * first_in_scale, through the same kernel (eps unused: cx=1, ce=0)
        self.o32.cfg_step(self.latents, self.latents, self.latents, 0.0, 1.0, 0.0, scheduler.first_in_scale(),
                          latents_out=self.latents, model_in=self.model_in)

    def capture(self):
        if not self.use_graph or self.graph is not None:
            return
        self.params.copy_(torch.tensor([500.0, 500.0, 0, 0, 1.0, 1.0, 0.0, 1.0]))
        if self.kv is None:
            self.kv = self.unet.text_kv(self.text)
        keep_l, keep_m = self.latents.clone(), self.model_in.clone()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):                 # warm-up: packs weights, fills the text K/V cache, sets func attrs
                self._body()
        torch.cuda.current_stream().wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._body()
        self.graph = g
        self.latents.copy_(keep_l)
        self.model_in.copy_(keep_m)

    def step(self, i: int):
        self.params.copy_(self.table[i], non_blocking=True)
        if self.graph is not None:
            self.graph.replay()
        else:
            self._body()


class AnimateDiffPipeline:
    """diffusers.AnimateDiffPipeline surface over the b200ad denoise step."""

    vae_scale_factor = 8

    def __init__(self, unet: UNetMotionModel, scheduler, vae: Optional[AutoencoderKLDecoder] = None,
                 torch_dtype=None, cfg_group=None):
        self.unet = unet
        self.scheduler = scheduler
        self.vae = vae
        self.dtype = torch_dtype or torch.float32
        self.device = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
        self.cfg_group = cfg_group              # 2-rank process group: CFG halves split across a GPU pair
        self.use_graph = os.environ.get("B200AD_GRAPH", "1") != "0"
        self._plans = {}
        self.last_latents = None

    # ---- loaders / toggles ---------------------------------------------------------------------
    @classmethod
    def from_pretrained(cls, name_or_path: str, motion_adapter: Optional[MotionAdapter] = None, torch_dtype=None,
                        seed: int = 0, mode: Optional[str] = None, with_vae: bool = True, **kw):
        """Local diffusers-layout directory if present; otherwise seeded random init of the same architecture
        (no network / no checkpoints on this box) -- logged loudly."""
        if not torch.cuda.is_available():
            raise RuntimeError("b200ad: a CUDA device is required (there is no CPU fallback)")
        mode = mode or os.environ.get("B200AD_MODE", "bf16")
        rng = torch.random.get_rng_state()
        crng = torch.cuda.get_rng_state()
        torch.manual_seed(seed)
        with torch.device("cuda"):
            unet = UNetMotionModel(UNetMotionConfig(), mode=mode)
            vae = AutoencoderKLDecoder() if with_vae else None
        torch.random.set_rng_state(rng)
        torch.cuda.set_rng_state(crng)
        sd = _load_local_state_dict(name_or_path, "unet")
        if sd is not None:
            unet.load_state_dict(sd, strict=False)
        else:
            _log(f"'{name_or_path}': no local checkpoint -> SD1.5 UNet + motion modules use seeded random init "
                 f"(seed {seed}); prompts map to synthetic text embeddings")
        if motion_adapter is not None and motion_adapter.state_dict_ is not None:
            unet.load_state_dict(motion_adapter.state_dict_, strict=False)
        if vae is not None:
            vsd = _load_local_state_dict(name_or_path, "vae")
            if vsd is not None:
                vae.load_state_dict(vsd, strict=False)
        sched = _sched.PNDMPlaceholder(_sched.SD15_SCHEDULER_CONFIG)
        return cls(unet, sched, vae, torch_dtype=torch_dtype)

    def to(self, *a, **kw):
        return self

    def enable_vae_slicing(self):
        if self.vae is not None:
            self.vae.use_slicing = True

    def disable_vae_slicing(self):
        if self.vae is not None:
            self.vae.use_slicing = False

    def enable_model_cpu_offload(self, *a, **kw):
        """No-op: 180 GB of HBM holds the whole pipeline (2.6 GB of bf16 weights)."""

    def maybe_free_model_hooks(self):
        pass

    def set_progress_bar_config(self, **kw):
        pass

    # ---- pieces of __call__ -------------------------------------------------------------------
    def encode_prompt(self, prompt, negative_prompt, do_cfg: bool):
        pe = synthetic_text_embeds(prompt)[None]
        ne = synthetic_text_embeds(negative_prompt or "")[None] if do_cfg else None
        return pe, ne

    def prepare_latents(self, num_frames, height, width, generator, latents=None):
        shape = (1, 4, num_frames, height // self.vae_scale_factor, width // self.vae_scale_factor)
        if latents is None:
            gdev = generator.device if generator is not None else torch.device("cpu")
            # diffusers randn_tensor: a CPU generator samples on the CPU in the pipeline dtype, then moves
            latents = torch.randn(shape, generator=generator, device=gdev, dtype=self.dtype)
        elif tuple(latents.shape) != shape:
            raise ValueError(f"latents shape {tuple(latents.shape)} != {shape}")
        return latents.to("cuda", non_blocking=True).float() * float(self.scheduler.init_noise_sigma)

    def _plan(self, B, F, h, w) -> DenoisePlan:
        key = (B, F, h, w, self.unet.mode, self.use_graph, self.cfg_group is not None)
        if key not in self._plans:
            self._plans[key] = DenoisePlan(self.unet, B, F, h, w, self.use_graph, self.cfg_group)
        return self._plans[key]

    @torch.no_grad()
    def denoise(self, latents, text, guidance_scale: float, num_inference_steps: int, callback=None):
        """The loop body of AnimateDiffPipeline.__call__ for our schedulers: latents
        [1,4,F,h,w] (already scaled by init_noise_sigma), text [B,77,768] with the uncond row first."""
        sch = self.scheduler
        sch.set_timesteps(num_inference_steps, device=None)
        B = text.shape[0]
        _, _, F, h, w = latents.shape
        plan = self._plan(B, F, h, w)
        if plan.split:
            text = text[torch.distributed.get_rank(self.cfg_group)][None]
        plan.load(latents, text, sch, guidance_scale)
        plan.capture()
        for i in range(num_inference_steps):
            plan.step(i)
            if callback is not None:
                callback(i, sch._timesteps_host[i], plan.latents)
        return plan.latents

    @torch.no_grad()
    def _denoise_generic(self, latents, text, guidance_scale, num_inference_steps):
        """Duck-typed loop for a foreign scheduler object (set_timesteps / scale_model_input / step)."""
        sch = self.scheduler
        sch.set_timesteps(num_inference_steps, device="cuda")
        do_cfg = text.shape[0] == 2
        for t in sch.timesteps:
            x_in = torch.cat([latents] * 2) if do_cfg else latents
            x_in = sch.scale_model_input(x_in, t)
            eps = self.unet(x_in, t, encoder_hidden_states=text).sample
            if do_cfg:
                u, c = eps.chunk(2)
                eps = u + guidance_scale * (c - u)
            latents = sch.step(eps, t, latents).prev_sample
        return latents

    @torch.no_grad()
    def decode_latents(self, latents):
        if self.vae is None:
            raise RuntimeError("pipeline was built without a VAE (with_vae=False); use output_type='latent'")
        z = (latents / self.vae.scaling_factor)[0].permute(1, 0, 2, 3)          # [F, 4, h, w]
        img = torch.cat([self.vae.decode(z[i:i + 16]) for i in range(0, z.shape[0], 16)])
        return img[None].permute(0, 2, 1, 3, 4).float()                          # [1, 3, F, H, W]

    @staticmethod
    def postprocess_video(video, output_type="pil"):
        v = (video / 2 + 0.5).clamp(0, 1)
        if output_type == "pt":
            return v.permute(0, 2, 1, 3, 4)
        arr = (v.permute(0, 2, 3, 4, 1) * 255).round().to(torch.uint8).cpu().numpy()   # [1, F, H, W, 3]
        if output_type == "np":
            return arr.astype(np.float32) / 255.0
        from PIL import Image
        return [[Image.fromarray(f) for f in vid] for vid in arr]

    # ---- the call ------------------------------------------------------------------------------
    @torch.no_grad()
    def __call__(self, prompt: Optional[str] = None, num_frames: int = 16, height: Optional[int] = None,
                 width: Optional[int] = None, num_inference_steps: int = 50, guidance_scale: float = 7.5,
                 negative_prompt: Optional[str] = None, generator=None, latents=None, prompt_embeds=None,
                 negative_prompt_embeds=None, output_type: str = "pil", return_dict: bool = True, eta: float = 0.0,
                 callback_on_step_end=None, **kw):
        height = height or 512
        width = width or 512
        if height % 8 or width % 8:
            raise ValueError(f"`height` and `width` have to be divisible by 8 but are {height} and {width}.")
        if isinstance(prompt, (list, tuple)):
            if len(prompt) != 1:
                raise NotImplementedError("batch size 1 (the reference's use) is implemented")
            prompt = prompt[0]
        if isinstance(negative_prompt, (list, tuple)):
            negative_prompt = negative_prompt[0]
        if isinstance(generator, (list, tuple)):
            generator = generator[0]
        if eta != 0.0:
            raise NotImplementedError("eta != 0 is not on the reference's path")
        do_cfg = guidance_scale > 1.0
        if prompt_embeds is None:
            if prompt is None:
                raise ValueError("Provide either `prompt` or `prompt_embeds`.")
            prompt_embeds, neg = self.encode_prompt(prompt, negative_prompt, do_cfg)
            if negative_prompt_embeds is None:
                negative_prompt_embeds = neg
        if do_cfg:
            if negative_prompt_embeds is None:
                negative_prompt_embeds = synthetic_text_embeds("")[None]
            text = torch.cat([negative_prompt_embeds, prompt_embeds])           # uncond row FIRST
        else:
            text = prompt_embeds
        text = text.to("cuda", non_blocking=True).float()
        fast = hasattr(self.scheduler, "step_coefficients")
        if not fast and not hasattr(self.scheduler, "set_timesteps"):
            raise TypeError("pipe.scheduler must be a scheduler object (assign DDIMScheduler.from_config(...))")
        lat = self.prepare_latents(num_frames, height, width, generator, latents)
        if fast:
            lat = self.denoise(lat, text, guidance_scale, num_inference_steps, callback_on_step_end)
        else:
            lat = self._denoise_generic(lat, text, guidance_scale, num_inference_steps)
        self.last_latents = lat
        if output_type == "latent":
            out = lat.to(self.dtype).clone()
        else:
            out = self.postprocess_video(self.decode_latents(lat), output_type)
        return AnimateDiffPipelineOutput(frames=out) if return_dict else (out,)

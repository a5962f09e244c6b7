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


_CKPT_NAMES = ("diffusion_pytorch_model.safetensors", "diffusion_pytorch_model.fp16.safetensors",
               "diffusion_pytorch_model.bin", "diffusion_pytorch_model.fp16.bin")


def _load_local_state_dict(path: str, sub: str):
    """State dict of a local diffusers-layout directory (``<path>/<sub>/diffusion_pytorch_model[.fp16].safetensors``
    or ``.bin``); None when the directory holds no checkpoint (this box: no network, no weights)."""
    d = os.path.join(path, sub) if sub else path
    for name in _CKPT_NAMES:
        f = os.path.join(d, name)
        if os.path.isfile(f):
            if name.endswith(".safetensors"):
                from safetensors.torch import load_file
                return load_file(f)
            return torch.load(f, map_location="cpu", weights_only=True)
    return None


_LEGACY_VAE_ATTN = {"query": "to_q", "key": "to_k", "value": "to_v", "proj_attn": "to_out.0"}


def _rename_legacy_vae_keys(sd: dict) -> dict:
    """SD1.5's vae checkpoint stores the mid-block attention under the deprecated AttentionBlock names
    (query/key/value/proj_attn); diffusers renames them on load, and so must we."""
    out = {}
    for k, v in sd.items():
        parts = k.split(".")
        if "attentions" in parts and len(parts) >= 2 and parts[-2] in _LEGACY_VAE_ATTN:
            parts[-2:-1] = _LEGACY_VAE_ATTN[parts[-2]].split(".")
            k = ".".join(parts)
            if v.dim() == 4:                        # very old checkpoints: 1x1 conv weights
                v = v[:, :, 0, 0]
        out[k] = v
    return out


def _load_checked(module, sd: dict, what: str, allow_missing=lambda k: False, allow_unexpected=lambda k: False):
    """load_state_dict(strict=False) that REFUSES silent partial loads: any parameter of `module` the checkpoint does
    not provide (other than the ones `allow_missing` names) or any checkpoint key the module does not know raises."""
    res = module.load_state_dict(sd, strict=False)
    missing = [k for k in res.missing_keys if not allow_missing(k)]
    unexpected = [k for k in res.unexpected_keys if not allow_unexpected(k)]
    if missing or unexpected:
        raise RuntimeError(
            f"{what}: checkpoint does not match the module tree -- {len(missing)} missing "
            f"(e.g. {missing[:3]}), {len(unexpected)} unexpected (e.g. {unexpected[:3]}); refusing to run with "
            "partially random weights")
    _log(f"{what}: loaded {len(sd) - len(res.unexpected_keys)} tensors"
         + (f" ({len(res.missing_keys)} left at their init by design)" if res.missing_keys else ""))


class _LocalClip:
    """CLIP ViT-L/14 text encoder + tokenizer from a LOCAL diffusers-layout directory (``text_encoder/``,
    ``tokenizer/``), as AnimateDiffPipeline loads them at /root/reference/experiments/05_grid_search_ablation.py:130-134.
    The encoder runs on the b200ad kernels (clip.py), once per prompt; the tokenizer is transformers'
    CLIPTokenizer (host-side string processing)."""

    def __init__(self, tokenizer, model):
        self.tokenizer, self.model = tokenizer, model

    @torch.no_grad()
    def __call__(self, prompt: str) -> torch.Tensor:
        ids = self.tokenizer(prompt, padding="max_length", max_length=self.tokenizer.model_max_length,
                             truncation=True, return_tensors="pt").input_ids
        return self.model(ids.cuda()).float()                              # [1, 77, 768]


def _load_text_encoder(path: str):
    te, tk = os.path.join(path, "text_encoder"), os.path.join(path, "tokenizer")
    if not (os.path.isdir(te) and os.path.isdir(tk)):
        _log(f"'{path}': no local text_encoder/ + tokenizer/ -> prompts map to synthetic text embeddings "
             "(seeded by the prompt string)")
        return None
    from transformers import CLIPTokenizer
    from .clip import CLIPTextEncoder
    tok = CLIPTokenizer.from_pretrained(tk, local_files_only=True)
    model = CLIPTextEncoder.from_local(te)
    _log(f"'{path}': CLIP text encoder loaded from the local directory (b200ad kernels)")
    return _LocalClip(tok, model)


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
        key = (type(scheduler).__name__, tuple(sorted((k, repr(v)) for k, v in dict(scheduler.config).items())), n,
               float(guidance_scale), self.B)
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

    def set_latents(self, latents, scheduler, i: int):
        """Replace the latents after step i (callback_on_step_end returned new ones): also refresh the next model input."""
        self.latents.copy_(latents.reshape(self.latents.shape).float())
        nxt = scheduler.step_coefficients(i)[2]
        self.o32.cfg_step(self.latents, self.latents, self.latents, 0.0, 1.0, 0.0, nxt,
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
        self.text_encoder = None                # _LocalClip when a local checkpoint directory provides one
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
        is_motion = lambda k: "motion_modules" in k
        sd = _load_local_state_dict(name_or_path, "unet")
        if sd is not None:
            # the SD1.5 UNet2DConditionModel checkpoint holds every non-motion parameter
            _load_checked(unet, sd, f"unet of '{name_or_path}'", allow_missing=is_motion)
        else:
            _log(f"'{name_or_path}': no local checkpoint -> SD1.5 UNet + motion modules use seeded random init "
                 f"(seed {seed})")
        if motion_adapter is not None and motion_adapter.state_dict_ is not None:
            # a MotionAdapter checkpoint holds exactly the motion-module subset
            _load_checked(unet, motion_adapter.state_dict_, f"motion adapter '{motion_adapter.source}'",
                          allow_missing=lambda k: not is_motion(k))
        if vae is not None:
            vsd = _load_local_state_dict(name_or_path, "vae")
            if vsd is not None:
                _load_checked(vae, _rename_legacy_vae_keys(vsd), f"vae of '{name_or_path}'",
                              allow_unexpected=lambda k: k.startswith(("encoder.", "quant_conv.")))
        sched = _sched.PNDMPlaceholder(_sched.SD15_SCHEDULER_CONFIG)
        pipe = cls(unet, sched, vae, torch_dtype=torch_dtype)
        pipe.text_encoder = _load_text_encoder(name_or_path)
        return pipe

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
        """[1,77,768] embeddings of the prompt (and of the negative prompt under CFG): the local CLIP text encoder
        when the checkpoint directory has one, else deterministic synthetic embeddings seeded by the string."""
        enc = self.text_encoder if self.text_encoder is not None else (lambda p: synthetic_text_embeds(p)[None])
        pe = enc(prompt)
        ne = enc(negative_prompt or "") if do_cfg else None
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
        if sch.num_inference_steps != num_inference_steps:
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
                # diffusers: callback_on_step_end(pipe, step_index, timestep, callback_kwargs) -> dict; a returned
                # "latents" replaces the loop's latents
                ret = callback(self, i, sch.timesteps[i], {"latents": plan.latents})
                if isinstance(ret, dict) and ret.get("latents") is not None and ret["latents"] is not plan.latents:
                    plan.set_latents(ret["latents"], sch, i)
        return plan.latents.clone()              # the plan's buffer is reused by the next call

    @torch.no_grad()
    def _denoise_generic(self, latents, text, guidance_scale, num_inference_steps, callback=None):
        """Duck-typed loop for a foreign scheduler object (set_timesteps / scale_model_input / step)."""
        sch = self.scheduler
        do_cfg = text.shape[0] == 2
        for i, t in enumerate(sch.timesteps):
            x_in = torch.cat([latents] * 2) if do_cfg else latents
            x_in = sch.scale_model_input(x_in, t)
            eps = self.unet(x_in, t, encoder_hidden_states=text).sample
            if do_cfg:
                u, c = eps.chunk(2)
                eps = u + guidance_scale * (c - u)
            latents = sch.step(eps, t, latents).prev_sample
            if callback is not None:
                ret = callback(self, i, t, {"latents": latents})
                if isinstance(ret, dict) and ret.get("latents") is not None:
                    latents = ret["latents"]
        return latents

    @torch.no_grad()
    def decode_latents(self, latents):
        """latents [1,4,F,h,w] -> video [1,3,F,8h,8w] fp32 (AutoencoderKL decode on the b200ad kernels, vae.py)."""
        if self.vae is None:
            raise RuntimeError("pipeline was built without a VAE (with_vae=False); use output_type='latent'")
        return self.vae.decode_video(latents.float())

    def postprocess_video(self, video, output_type="pil"):
        """diffusers VideoProcessor.postprocess_video: (x/2+0.5).clamp(0,1) -> 'pt' [1,F,3,H,W] float, 'np'
        [1,F,H,W,3] float, 'pil' list of lists of PIL images.  The uint8 conversion is b200ad_video_to_u8."""
        u8 = get_ops("fp32").video_to_u8(video[0].float().contiguous())              # [F, H, W, 3] uint8 on the device
        if output_type == "pt":
            return (u8.permute(0, 3, 1, 2).float() / 255.0)[None]
        arr = u8.cpu().numpy()
        if output_type == "np":
            return arr.astype(np.float32)[None] / 255.0
        from PIL import Image
        return [[Image.fromarray(f) for f in arr]]

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
        # diffusers order: set_timesteps BEFORE prepare_latents -- init_noise_sigma of EulerDiscreteScheduler
        # ("leading" spacing) depends on the sigma table of THIS call's step count
        self.scheduler.set_timesteps(num_inference_steps, **({} if fast else {"device": "cuda"}))
        lat = self.prepare_latents(num_frames, height, width, generator, latents)
        if fast:
            lat = self.denoise(lat, text, guidance_scale, num_inference_steps, callback_on_step_end)
        else:
            lat = self._denoise_generic(lat, text, guidance_scale, num_inference_steps, callback_on_step_end)
        self.last_latents = lat
        if output_type == "latent":
            out = lat.to(self.dtype).clone()
        else:
            out = self.postprocess_video(self.decode_latents(lat), output_type)
        return AnimateDiffPipelineOutput(frames=out) if return_dict else (out,)

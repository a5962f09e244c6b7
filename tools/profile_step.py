#!/usr/bin/env python
"""One eager (no CUDA graph) denoise step of the bench workload, for ncu:
    python tools/profile_step.py [--steps S]        # S eager steps after one warm-up step (default 1)
Step k's launches are [k*L, (k+1)*L) with L printed at the end (ncu: -s L -c L captures exactly one warm step)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["B200AD_GRAPH"] = "0"

import torch  # noqa: E402
import b200ad  # noqa: E402,F401
from b200ad import lib  # noqa: E402
from b200ad.pipeline import AnimateDiffPipeline  # noqa: E402
from b200ad.schedulers import DDIMScheduler  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--frames", type=int, default=16)
ap.add_argument("--latent", type=int, default=64)
a = ap.parse_args()
pipe = AnimateDiffPipeline.from_pretrained("runwayml/stable-diffusion-v1-5", seed=0, with_vae=False)
pipe.scheduler = DDIMScheduler.from_config(pipe.scheduler.config, beta_schedule="linear", steps_offset=1,
                                           clip_sample=False)
pipe.scheduler.set_timesteps(25)
pipe.unet.prepack()
plan = pipe._plan(2, a.frames, a.latent, a.latent)
lat = torch.randn(1, 4, a.frames, a.latent, a.latent, generator=torch.Generator().manual_seed(42))
txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43))
plan.load(lat.cuda(), txt.cuda(), pipe.scheduler, 7.5)
torch.cuda.synchronize()
n0 = lib.launch_count()
plan.step(0)
torch.cuda.synchronize()
L = lib.launch_count() - n0
for i in range(a.steps):
    plan.step(1 + i)
torch.cuda.synchronize()
print(f"launches_before_first_step={n0} launches_per_step={L}")

#!/usr/bin/env python
"""One UNet forward per kernel mode on a reduced-width model, meant to run under compute-sanitizer:

    PYTORCH_NO_CUDA_MEMORY_CACHING=1 compute-sanitizer --tool initcheck python tools/sanitize_forward.py
    compute-sanitizer --tool memcheck  python tools/sanitize_forward.py
    compute-sanitizer --tool racecheck python tools/sanitize_forward.py --latent 16

With the caching allocator off every torch.empty() is a fresh cudaMalloc, so initcheck reports any kernel that reads
bytes no kernel of THIS forward has written (pad columns, statistics slots, tile tails ...)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import b200ad  # noqa: E402,F401
from b200ad.unet import UNetMotionConfig, UNetMotionModel  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--latent", type=int, default=64)
ap.add_argument("--frames", type=int, default=16)
ap.add_argument("--modes", default="bf16,fp32")
ap.add_argument("--vae", action="store_true")
a = ap.parse_args()
torch.manual_seed(0)
m = UNetMotionModel(UNetMotionConfig(block_out_channels=(64, 128, 256, 256))).cuda()
lat = torch.randn(2, 4, a.frames, a.latent, a.latent, device="cuda")
txt = torch.randn(2, 77, 768, device="cuda")
for mode in a.modes.split(","):
    m.set_mode(mode)
    out = m(lat, 500, encoder_hidden_states=txt).sample
    torch.cuda.synchronize()
    print(mode, "forward done, finite:", bool(torch.isfinite(out).all()), float(out.abs().mean()), flush=True)
if a.vae:
    from b200ad.vae import AutoencoderKLDecoder
    v = AutoencoderKLDecoder((128, 128, 64, 64)).cuda()
    img = v.decode_video(torch.randn(1, 4, 2, 16, 16, device="cuda"))
    torch.cuda.synchronize()
    print("vae decode done", tuple(img.shape), flush=True)

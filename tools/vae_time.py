#!/usr/bin/env python
"""VAE decode + uint8 postprocess of one 16-frame 512x512 video (latents [1,4,16,64,64]), CUDA events, 5 runs after
2 warm-ups:  python tools/vae_time.py   (B200AD_UPCONV=0 for the direct upsample + conv form)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import b200ad  # noqa: E402,F401
from b200ad.vae import AutoencoderKLDecoder  # noqa: E402

torch.manual_seed(0)
vae = AutoencoderKLDecoder().cuda()
lat = torch.randn(1, 4, 16, 64, 64, device="cuda")
for _ in range(2):
    vae.decode_video_u8(lat)
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); vae.decode_video_u8(lat); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print(f"decode_video_u8 16f 512^2: {min(ts):.1f} ms best, {sum(ts) / len(ts):.1f} ms mean "
      f"(B200AD_UPCONV={os.environ.get('B200AD_UPCONV', '1')})")

"""AutoencoderKL *decoder* of SD1.5, used once per video by ``decode_latents``
(/root/reference/experiments/05_grid_search_ablation.py:169 consumes its frames).

Scope note: the VAE is row f1 ("next") of SURVEY.md section 8 -- it is not part of the denoise step the sm_100a
kernels cover.  Until that row is built it runs as plain PyTorch modules (state-dict keys follow diffusers'
``vae.decoder.*`` / ``vae.post_quant_conv.*``), on whatever device the pipeline lives on.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class _VaeResnet(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.norm1 = nn.GroupNorm(32, cin, eps=1e-6)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.norm2 = nn.GroupNorm(32, cout, eps=1e-6)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x):
        h = self.conv1(F.silu(self.norm1(x)))
        h = self.conv2(F.silu(self.norm2(h)))
        return (x if self.conv_shortcut is None else self.conv_shortcut(x)) + h


class _VaeAttention(nn.Module):
    """Single-head attention over h*w tokens, d = 512, q/k/v/out with bias."""

    def __init__(self, c):
        super().__init__()
        self.group_norm = nn.GroupNorm(32, c, eps=1e-6)
        self.to_q, self.to_k, self.to_v = nn.Linear(c, c), nn.Linear(c, c), nn.Linear(c, c)
        self.to_out = nn.ModuleList([nn.Linear(c, c), nn.Dropout(0.0)])

    def forward(self, x):
        n, c, h, w = x.shape
        t = self.group_norm(x).view(n, c, h * w).transpose(1, 2)
        q, k, v = self.to_q(t)[:, None], self.to_k(t)[:, None], self.to_v(t)[:, None]
        o = F.scaled_dot_product_attention(q, k, v)[:, 0]
        o = self.to_out[0](o).transpose(1, 2).reshape(n, c, h, w)
        return x + o


class _VaeMid(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.resnets = nn.ModuleList([_VaeResnet(c, c), _VaeResnet(c, c)])
        self.attentions = nn.ModuleList([_VaeAttention(c)])

    def forward(self, x):
        return self.resnets[1](self.attentions[0](self.resnets[0](x)))


class _VaeUpsample(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class _VaeUpBlock(nn.Module):
    def __init__(self, cin, cout, up):
        super().__init__()
        self.resnets = nn.ModuleList([_VaeResnet(cin if i == 0 else cout, cout) for i in range(3)])
        self.upsamplers = nn.ModuleList([_VaeUpsample(cout)]) if up else None

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        return x if self.upsamplers is None else self.upsamplers[0](x)


class _Decoder(nn.Module):
    def __init__(self, chans=(512, 512, 256, 128)):
        super().__init__()
        self.conv_in = nn.Conv2d(4, chans[0], 3, padding=1)
        self.mid_block = _VaeMid(chans[0])
        self.up_blocks = nn.ModuleList()
        prev = chans[0]
        for i, c in enumerate(chans):
            self.up_blocks.append(_VaeUpBlock(prev, c, up=i < len(chans) - 1))
            prev = c
        self.conv_norm_out = nn.GroupNorm(32, prev, eps=1e-6)
        self.conv_out = nn.Conv2d(prev, 3, 3, padding=1)

    def forward(self, z):
        x = self.mid_block(self.conv_in(z))
        for b in self.up_blocks:
            x = b(x)
        return self.conv_out(F.silu(self.conv_norm_out(x)))


class AutoencoderKLDecoder(nn.Module):
    scaling_factor = 0.18215

    def __init__(self, chans=(512, 512, 256, 128)):
        super().__init__()
        self.post_quant_conv = nn.Conv2d(4, 4, 1)
        self.decoder = _Decoder(chans)
        self.use_slicing = False
        for p in self.parameters():
            p.requires_grad_(False)

    @torch.no_grad()
    def decode(self, z):
        if self.use_slicing and z.shape[0] > 1:
            return torch.cat([self.decoder(self.post_quant_conv(s)) for s in z.split(1)])
        return self.decoder(self.post_quant_conv(z))

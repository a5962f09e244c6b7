"""Oracle: fp32 CPU restatement of the SD1.5 AutoencoderKL *decoder* + diffusers' video postprocess.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED: diffusers is not importable on this image
and the reference holds no decoded-tensor fixture; the restatement follows SURVEY.md Appendix A.2.8 (diffusers
``AutoencoderKL`` as SD1.5 configures it) and is consumed by the reference at
/root/reference/experiments/05_grid_search_ablation.py:169 (``output.frames[0]``) and :178-179 (``frame.save``).
The PSNR used by the gate is the reference's own formula,
/root/reference/experiments/06_measure_grid_search.py:209-218 (``10*log10(1/mse)`` on [0,1] images).

Written as the textbook sequence of torch ops in NCHW, one function per diffusers module; state-dict keys follow
diffusers (``decoder.up_blocks.i.resnets.j.conv1.weight`` ...), so the same weights load into the product decoder.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

SCALING_FACTOR = 0.18215          # SD1.5 vae/config.json "scaling_factor"


class VaeResnetBlock(nn.Module):
    """diffusers ResnetBlock2D with temb_channels=None, groups 32, eps 1e-6, output_scale_factor 1."""

    def __init__(self, cin: int, cout: int, groups: int = 32):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=1e-6)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.norm2 = nn.GroupNorm(groups, cout, eps=1e-6)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x):
        h = self.norm1(x)
        h = F.silu(h)
        h = self.conv1(h)
        h = self.norm2(h)
        h = F.silu(h)
        h = self.conv2(h)
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return x + h


class VaeMidAttention(nn.Module):
    """diffusers Attention(heads=1, dim_head=512, bias=True, norm_num_groups=32, residual_connection=True,
    _from_deprecated_attn_block): GroupNorm -> q,k,v Linear(+bias) -> softmax(q k^T / sqrt(512)) v -> Linear -> + x."""

    def __init__(self, c: int, groups: int = 32):
        super().__init__()
        self.group_norm = nn.GroupNorm(groups, c, eps=1e-6)
        self.to_q = nn.Linear(c, c)
        self.to_k = nn.Linear(c, c)
        self.to_v = nn.Linear(c, c)
        self.to_out = nn.ModuleList([nn.Linear(c, c), nn.Dropout(0.0)])

    def forward(self, x):
        n, c, h, w = x.shape
        tokens = self.group_norm(x).reshape(n, c, h * w).transpose(1, 2)          # [n, hw, c]
        q, k, v = self.to_q(tokens), self.to_k(tokens), self.to_v(tokens)
        scores = torch.bmm(q, k.transpose(1, 2)) * (1.0 / math.sqrt(c))
        probs = scores.softmax(dim=-1)
        o = torch.bmm(probs, v)
        o = self.to_out[0](o)
        return x + o.transpose(1, 2).reshape(n, c, h, w)


class VaeMidBlock(nn.Module):
    def __init__(self, c: int, groups: int = 32):
        super().__init__()
        self.attentions = nn.ModuleList([VaeMidAttention(c, groups)])
        self.resnets = nn.ModuleList([VaeResnetBlock(c, c, groups), VaeResnetBlock(c, c, groups)])

    def forward(self, x):
        x = self.resnets[0](x)
        x = self.attentions[0](x)
        return self.resnets[1](x)


class VaeUpsample(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class VaeUpDecoderBlock(nn.Module):
    def __init__(self, cin: int, cout: int, add_upsample: bool, groups: int = 32, layers: int = 3):
        super().__init__()
        self.resnets = nn.ModuleList([VaeResnetBlock(cin if j == 0 else cout, cout, groups) for j in range(layers)])
        self.upsamplers = nn.ModuleList([VaeUpsample(cout)]) if add_upsample else None

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class VaeDecoder(nn.Module):
    """diffusers Decoder: block_out_channels (128,256,512,512) walked in reverse, layers_per_block 2 (+1)."""

    def __init__(self, up_channels: Sequence[int] = (512, 512, 256, 128), groups: int = 32):
        super().__init__()
        self.conv_in = nn.Conv2d(4, up_channels[0], 3, padding=1)
        self.mid_block = VaeMidBlock(up_channels[0], groups)
        blocks, prev = [], up_channels[0]
        for i, c in enumerate(up_channels):
            blocks.append(VaeUpDecoderBlock(prev, c, add_upsample=i != len(up_channels) - 1, groups=groups))
            prev = c
        self.up_blocks = nn.ModuleList(blocks)
        self.conv_norm_out = nn.GroupNorm(groups, prev, eps=1e-6)
        self.conv_out = nn.Conv2d(prev, 3, 3, padding=1)

    def forward(self, z):
        x = self.conv_in(z)
        x = self.mid_block(x)
        for blk in self.up_blocks:
            x = blk(x)
        x = self.conv_norm_out(x)
        x = F.silu(x)
        return self.conv_out(x)


class OracleVaeDecoder(nn.Module):
    """AutoencoderKL.decode: post_quant_conv (1x1, 4->4) then the decoder."""

    def __init__(self, up_channels: Sequence[int] = (512, 512, 256, 128), groups: int = 32):
        super().__init__()
        self.post_quant_conv = nn.Conv2d(4, 4, 1)
        self.decoder = VaeDecoder(up_channels, groups)

    @torch.no_grad()
    def decode(self, z):
        return self.decoder(self.post_quant_conv(z))


def make_oracle_vae(seed: int = 1, up_channels: Sequence[int] = (512, 512, 256, 128), groups: int = 32,
                    round_to: Optional[torch.dtype] = torch.bfloat16) -> OracleVaeDecoder:
    """Default torch init under ``seed``, parameters rounded once to the fast path's operand dtype and kept in fp32
    storage (the same "identical weights" convention as make_oracle_unet)."""
    state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    m = OracleVaeDecoder(up_channels, groups).eval()
    torch.random.set_rng_state(state)
    with torch.no_grad():
        for p in m.parameters():
            if round_to is not None:
                p.copy_(p.to(round_to).float())
            p.requires_grad_(False)
    return m


@torch.no_grad()
def decode_latents(vae: OracleVaeDecoder, latents: torch.Tensor) -> torch.Tensor:
    """AnimateDiffPipeline.decode_latents: latents [1,4,F,h,w] -> video [1,3,F,8h,8w] (float32, roughly [-1,1])."""
    z = latents / SCALING_FACTOR
    b, c, f, h, w = z.shape
    z = z.permute(0, 2, 1, 3, 4).reshape(b * f, c, h, w)
    img = vae.decode(z)
    return img[None, :].reshape((b, f, -1) + img.shape[2:]).permute(0, 2, 1, 3, 4).float()


def postprocess_video(video: torch.Tensor) -> torch.Tensor:
    """VideoProcessor.postprocess_video up to the uint8 frames: (x/2+0.5).clamp(0,1) -> [F,H,W,3] uint8 (round)."""
    v = (video / 2 + 0.5).clamp(0, 1)
    return (v[0].permute(1, 2, 3, 0) * 255).round().to(torch.uint8)


def psnr(a: torch.Tensor, b: torch.Tensor) -> float:
    """The reference's PSNR on [0,1] images (06_measure_grid_search.py:209-218)."""
    mse = F.mse_loss(a.float(), b.float()).item()
    if mse < 1e-10:
        return 100.0
    return 10 * math.log10(1.0 / mse)

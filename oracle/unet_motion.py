"""Oracle: fp32 CPU restatement of diffusers' UNetMotionModel (SD1.5 UNet + motion adapter v1.5-2).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED beyond the structural
known-answers of /root/reference/docs/02_video_diffusion_architecture.md:86-91 (1312.7 M params,
639 temporal modules, 32 spatial attentions; predicates at
/root/reference/experiments/02_architecture_inspection.py:51-60).

Module / state-dict names follow diffusers so that the same state_dict
loads into the product's module tree.  Everything here is written as the textbook sequence of
torch ops, NCHW, no fusion: it is the definition the CUDA path is compared against.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


# ----------------------------------------------------------------------------- configuration
@dataclass
class UNetMotionConfig:
    """SURVEY.md Appendix A.1 (SD1.5 UNet config + motion adapter v1-5-2 config)."""
    in_channels: int = 4
    out_channels: int = 4
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    layers_per_block: int = 2
    num_attention_heads: int = 8          # SD1.5 "attention_head_dim=8" means 8 heads
    cross_attention_dim: int = 768
    norm_num_groups: int = 32
    norm_eps: float = 1e-5
    motion_num_attention_heads: int = 8
    motion_max_seq_length: int = 32
    motion_norm_num_groups: int = 32
    time_embed_dim_mult: int = 4
    cross_attn_levels: Tuple[bool, ...] = (True, True, True, False)

    @property
    def time_embed_dim(self) -> int:
        return self.block_out_channels[0] * self.time_embed_dim_mult


# ----------------------------------------------------------------------------- attention pieces
class Attention(nn.Module):
    """diffusers Attention + AttnProcessor2_0.  No q/k/v bias, bias on to_out[0]."""

    def __init__(self, query_dim: int, heads: int, cross_attention_dim: Optional[int] = None):
        super().__init__()
        self.heads = heads
        self.inner_dim = query_dim
        kv_dim = query_dim if cross_attention_dim is None else cross_attention_dim
        self.to_q = nn.Linear(query_dim, query_dim, bias=False)
        self.to_k = nn.Linear(kv_dim, query_dim, bias=False)
        self.to_v = nn.Linear(kv_dim, query_dim, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(query_dim, query_dim, bias=True), nn.Dropout(0.0)])

    def forward(self, hidden_states, encoder_hidden_states=None):
        ctx = hidden_states if encoder_hidden_states is None else encoder_hidden_states
        q, k, v = self.to_q(hidden_states), self.to_k(ctx), self.to_v(ctx)
        bt, lq, c = q.shape
        d = c // self.heads
        q = q.view(bt, lq, self.heads, d).transpose(1, 2)
        k = k.view(bt, -1, self.heads, d).transpose(1, 2)
        v = v.view(bt, -1, self.heads, d).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=None, dropout_p=0.0, is_causal=False)
        o = o.transpose(1, 2).reshape(bt, lq, c)
        return self.to_out[1](self.to_out[0](o))


class GEGLU(nn.Module):
    def __init__(self, dim_in: int, dim_out: int):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)

    def forward(self, x):
        h, g = self.proj(x).chunk(2, dim=-1)
        return h * F.gelu(g)          # exact erf GELU


class FeedForward(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * 4), nn.Dropout(0.0), nn.Linear(dim * 4, dim)])

    def forward(self, x):
        for m in self.net:
            x = m(x)
        return x


class SinusoidalPositionalEmbedding(nn.Module):
    """pe[0,:,0::2]=sin(pos*div), pe[0,:,1::2]=cos(pos*div); persistent buffer."""

    def __init__(self, embed_dim: int, max_seq_length: int = 32):
        super().__init__()
        position = torch.arange(max_seq_length).unsqueeze(1)
        div_term = torch.exp(torch.arange(0, embed_dim, 2) * (-math.log(10000.0) / embed_dim))
        pe = torch.zeros(1, max_seq_length, embed_dim)
        pe[0, :, 0::2] = torch.sin(position * div_term)
        pe[0, :, 1::2] = torch.cos(position * div_term)
        self.register_buffer("pe", pe)

    def forward(self, x):
        return x + self.pe[:, : x.shape[1]]


class BasicTransformerBlock(nn.Module):
    """Spatial (cross_attention_dim=768, no PE) or temporal (double self-attention, sinusoidal PE)."""

    def __init__(self, dim: int, heads: int, cross_attention_dim: Optional[int], temporal: bool, max_seq: int = 32):
        super().__init__()
        self.temporal = temporal
        self.pos_embed = SinusoidalPositionalEmbedding(dim, max_seq) if temporal else None
        self.norm1 = nn.LayerNorm(dim, eps=1e-5)
        self.attn1 = Attention(dim, heads)
        self.norm2 = nn.LayerNorm(dim, eps=1e-5)
        self.attn2 = Attention(dim, heads, cross_attention_dim=None if temporal else cross_attention_dim)
        self.norm3 = nn.LayerNorm(dim, eps=1e-5)
        self.ff = FeedForward(dim)

    def forward(self, x, encoder_hidden_states=None):
        n = self.norm1(x)
        if self.pos_embed is not None:
            n = self.pos_embed(n)
        x = x + self.attn1(n)
        n = self.norm2(x)
        if self.pos_embed is not None:
            n = self.pos_embed(n)
        x = x + self.attn2(n, None if self.temporal else encoder_hidden_states)
        x = x + self.ff(self.norm3(x))
        return x


class Transformer2DModel(nn.Module):
    """SURVEY.md A.2.2 (use_linear_projection=False: 1x1 conv proj in/out, GN eps 1e-6)."""

    def __init__(self, channels: int, heads: int, cross_attention_dim: int, groups: int = 32):
        super().__init__()
        self.norm = nn.GroupNorm(groups, channels, eps=1e-6)
        self.proj_in = nn.Conv2d(channels, channels, 1)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(channels, heads, cross_attention_dim, temporal=False)])
        self.proj_out = nn.Conv2d(channels, channels, 1)

    def forward(self, x, encoder_hidden_states):
        n, c, h, w = x.shape
        r = x
        x = self.proj_in(self.norm(x))
        x = x.permute(0, 2, 3, 1).reshape(n, h * w, c)
        for blk in self.transformer_blocks:
            x = blk(x, encoder_hidden_states)
        x = x.reshape(n, h, w, c).permute(0, 3, 1, 2)
        return self.proj_out(x) + r


class AnimateDiffTransformer3D(nn.Module):
    """Motion module: GroupNorm statistics span the frame axis."""

    def __init__(self, channels: int, heads: int, groups: int = 32, max_seq: int = 32):
        super().__init__()
        self.norm = nn.GroupNorm(groups, channels, eps=1e-6)
        self.proj_in = nn.Linear(channels, channels)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(channels, heads, None, temporal=True, max_seq=max_seq)])
        self.proj_out = nn.Linear(channels, channels)

    def forward(self, x, num_frames: int = 1):
        bf, c, h, w = x.shape
        b = bf // num_frames
        r = x
        x = x[None, :].reshape(b, num_frames, c, h, w).permute(0, 2, 1, 3, 4)
        x = self.norm(x)
        x = x.permute(0, 3, 4, 2, 1).reshape(b * h * w, num_frames, c)
        x = self.proj_in(x)
        for blk in self.transformer_blocks:
            x = blk(x)
        x = self.proj_out(x)
        x = x[None, None, :].reshape(b, h, w, num_frames, c).permute(0, 3, 4, 1, 2).reshape(bf, c, h, w)
        return x + r


# ----------------------------------------------------------------------------- conv pieces
class ResnetBlock2D(nn.Module):
    """SURVEY.md A.2.1."""

    def __init__(self, cin: int, cout: int, temb_dim: int, groups: int = 32, eps: float = 1e-5):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb_dim, cout)
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.dropout = nn.Dropout(0.0)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.nonlinearity = nn.SiLU()
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x, temb):
        h = self.conv1(self.nonlinearity(self.norm1(x)))
        h = h + self.time_emb_proj(self.nonlinearity(temb))[:, :, None, None]
        h = self.conv2(self.dropout(self.nonlinearity(self.norm2(h))))
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return x + h


class Downsample2D(nn.Module):
    def __init__(self, channels: int):
        super().__init__()
        self.conv = nn.Conv2d(channels, channels, 3, stride=2, padding=1)

    def forward(self, x):
        return self.conv(x)


class Upsample2D(nn.Module):
    def __init__(self, channels: int):
        super().__init__()
        self.conv = nn.Conv2d(channels, channels, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


# ----------------------------------------------------------------------------- blocks
class DownBlockMotion(nn.Module):
    def __init__(self, cfg, cin, cout, cross: bool, add_downsample: bool):
        super().__init__()
        td = cfg.time_embed_dim
        self.resnets = nn.ModuleList()
        self.motion_modules = nn.ModuleList()
        if cross:
            self.attentions = nn.ModuleList()
        self.has_cross = cross
        for i in range(cfg.layers_per_block):
            self.resnets.append(ResnetBlock2D(cin if i == 0 else cout, cout, td, cfg.norm_num_groups, cfg.norm_eps))
            if cross:
                self.attentions.append(Transformer2DModel(cout, cfg.num_attention_heads, cfg.cross_attention_dim,
                                                          cfg.norm_num_groups))
            self.motion_modules.append(AnimateDiffTransformer3D(cout, cfg.motion_num_attention_heads,
                                                                cfg.motion_norm_num_groups, cfg.motion_max_seq_length))
        self.downsamplers = nn.ModuleList([Downsample2D(cout)]) if add_downsample else None

    def forward(self, x, temb, ctx, num_frames):
        skips = []
        for i, res in enumerate(self.resnets):
            x = res(x, temb)
            if self.has_cross:
                x = self.attentions[i](x, ctx)
            x = self.motion_modules[i](x, num_frames=num_frames)
            skips.append(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x)
            skips.append(x)
        return x, skips


class CrossAttnDownBlockMotion(DownBlockMotion):
    pass


class UNetMidBlockCrossAttnMotion(nn.Module):
    def __init__(self, cfg, c):
        super().__init__()
        td = cfg.time_embed_dim
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, td, cfg.norm_num_groups, cfg.norm_eps),
                                      ResnetBlock2D(c, c, td, cfg.norm_num_groups, cfg.norm_eps)])
        self.attentions = nn.ModuleList([Transformer2DModel(c, cfg.num_attention_heads, cfg.cross_attention_dim,
                                                            cfg.norm_num_groups)])
        self.motion_modules = nn.ModuleList([AnimateDiffTransformer3D(c, cfg.motion_num_attention_heads,
                                                                      cfg.motion_norm_num_groups,
                                                                      cfg.motion_max_seq_length)])

    def forward(self, x, temb, ctx, num_frames):
        x = self.resnets[0](x, temb)
        x = self.attentions[0](x, ctx)
        x = self.motion_modules[0](x, num_frames=num_frames)
        return self.resnets[1](x, temb)


class UpBlockMotion(nn.Module):
    def __init__(self, cfg, prev_c, out_c, skip_cs: List[int], cross: bool, add_upsample: bool):
        super().__init__()
        td = cfg.time_embed_dim
        self.resnets = nn.ModuleList()
        self.motion_modules = nn.ModuleList()
        if cross:
            self.attentions = nn.ModuleList()
        self.has_cross = cross
        for i, sc in enumerate(skip_cs):
            cin = (prev_c if i == 0 else out_c) + sc
            self.resnets.append(ResnetBlock2D(cin, out_c, td, cfg.norm_num_groups, cfg.norm_eps))
            if cross:
                self.attentions.append(Transformer2DModel(out_c, cfg.num_attention_heads, cfg.cross_attention_dim,
                                                          cfg.norm_num_groups))
            self.motion_modules.append(AnimateDiffTransformer3D(out_c, cfg.motion_num_attention_heads,
                                                                cfg.motion_norm_num_groups, cfg.motion_max_seq_length))
        self.upsamplers = nn.ModuleList([Upsample2D(out_c)]) if add_upsample else None

    def forward(self, x, skips, temb, ctx, num_frames):
        for i, res in enumerate(self.resnets):
            x = torch.cat([x, skips.pop()], dim=1)
            x = res(x, temb)
            if self.has_cross:
                x = self.attentions[i](x, ctx)
            x = self.motion_modules[i](x, num_frames=num_frames)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class CrossAttnUpBlockMotion(UpBlockMotion):
    pass


class Timesteps(nn.Module):
    """Sinusoidal timestep projection, flip_sin_to_cos=True, freq_shift=0."""

    def __init__(self, dim: int):
        super().__init__()
        self.dim = dim

    def forward(self, t):
        half = self.dim // 2
        freqs = torch.exp(-math.log(10000.0) * torch.arange(half, dtype=torch.float32) / half)
        e = t[:, None].float() * freqs[None]
        return torch.cat([torch.cos(e), torch.sin(e)], dim=-1)


class TimestepEmbedding(nn.Module):
    def __init__(self, cin: int, dim: int):
        super().__init__()
        self.linear_1 = nn.Linear(cin, dim)
        self.act = nn.SiLU()
        self.linear_2 = nn.Linear(dim, dim)

    def forward(self, x):
        return self.linear_2(self.act(self.linear_1(x)))


@dataclass
class UNetMotionOutput:
    sample: torch.Tensor


class UNetMotionModel(nn.Module):
    """SURVEY.md 3.3 / A.2.7.  forward(sample[B,4,F,h,w], timestep, encoder_hidden_states)."""

    def __init__(self, cfg: Optional[UNetMotionConfig] = None):
        super().__init__()
        cfg = cfg or UNetMotionConfig()
        self.cfg = cfg
        boc = cfg.block_out_channels
        self.conv_in = nn.Conv2d(cfg.in_channels, boc[0], 3, padding=1)
        self.time_proj = Timesteps(boc[0])
        self.time_embedding = TimestepEmbedding(boc[0], cfg.time_embed_dim)

        self.down_blocks = nn.ModuleList()
        skip_channels = [boc[0]]
        c = boc[0]
        for i, oc in enumerate(boc):
            last = i == len(boc) - 1
            cls = CrossAttnDownBlockMotion if cfg.cross_attn_levels[i] else DownBlockMotion
            self.down_blocks.append(cls(cfg, c, oc, cfg.cross_attn_levels[i], add_downsample=not last))
            skip_channels += [oc] * cfg.layers_per_block + ([oc] if not last else [])
            c = oc
        self.mid_block = UNetMidBlockCrossAttnMotion(cfg, boc[-1])

        self.up_blocks = nn.ModuleList()
        rev = list(reversed(boc))
        rev_cross = list(reversed(cfg.cross_attn_levels))
        prev = boc[-1]
        for i, oc in enumerate(rev):
            last = i == len(rev) - 1
            n = cfg.layers_per_block + 1
            scs = [skip_channels.pop() for _ in range(n)]
            cls = CrossAttnUpBlockMotion if rev_cross[i] else UpBlockMotion
            self.up_blocks.append(cls(cfg, prev, oc, scs, rev_cross[i], add_upsample=not last))
            prev = oc
        self.conv_norm_out = nn.GroupNorm(cfg.norm_num_groups, boc[0], eps=cfg.norm_eps)
        self.conv_act = nn.SiLU()
        self.conv_out = nn.Conv2d(boc[0], cfg.out_channels, 3, padding=1)

    def forward(self, sample, timestep, encoder_hidden_states):
        b, _, f, h, w = sample.shape
        t = timestep
        if not torch.is_tensor(t):
            t = torch.tensor([t], dtype=torch.float32 if isinstance(t, float) else torch.int64)
        elif t.dim() == 0:
            t = t[None]
        t = t.expand(b)
        emb = self.time_embedding(self.time_proj(t).to(sample.dtype))
        emb = emb.repeat_interleave(f, dim=0)
        ctx = encoder_hidden_states
        if ctx.shape[0] == b:                       # diffusers <=0.30 repeats inside the UNet
            ctx = ctx.repeat_interleave(f, dim=0)
        x = sample.permute(0, 2, 1, 3, 4).reshape(b * f, -1, h, w)
        x = self.conv_in(x)
        skips = [x]
        for blk in self.down_blocks:
            x, s = blk(x, emb, ctx, f)
            skips += s
        x = self.mid_block(x, emb, ctx, f)
        for blk in self.up_blocks:
            x = blk(x, skips, emb, ctx, f)
        x = self.conv_out(self.conv_act(self.conv_norm_out(x)))
        x = x[None, :].reshape(b, f, -1, h, w).permute(0, 2, 1, 3, 4)
        return UNetMotionOutput(sample=x)


# ----------------------------------------------------------------------------- shared-weights helper
def make_oracle_unet(seed: int = 0, cfg: Optional[UNetMotionConfig] = None,
                     round_to: Optional[torch.dtype] = torch.bfloat16) -> UNetMotionModel:
    """Random-init (torch default init under manual_seed) UNet whose parameters are rounded once to the
    fast path's operand dtype and kept in fp32 storage -- "the identical weights" of SURVEY.md 8(c)."""
    gen_state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    m = UNetMotionModel(cfg).eval()
    torch.random.set_rng_state(gen_state)
    if round_to is not None:
        with torch.no_grad():
            for p in m.parameters():
                p.copy_(p.to(round_to).float())
    for p in m.parameters():
        p.requires_grad_(False)
    return m

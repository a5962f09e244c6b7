"""Oracle: fp32 CPU restatement of the CLIP ViT-L/14 *text* encoder (``transformers.CLIPTextModel``) that
AnimateDiffPipeline loads as ``text_encoder`` (/root/reference/experiments/05_grid_search_ablation.py:130-134) and
runs once per prompt in ``encode_prompt`` -- SURVEY.md section 8 row f2.

TEST INFRASTRUCTURE ONLY.  PARITY PINNED: unlike diffusers, ``transformers`` (5.5) IS importable on this image, so
tests/test_oracle_cpu.py checks this restatement against the library's own CLIPTextModel on identical random-init
weights (no checkpoint exists offline).  State-dict keys follow transformers (``text_model.embeddings.token_embedding.
weight`` ...), so the same weights load into the library model, this oracle and the product encoder.

Architecture (openai/clip-vit-large-patch14 text config, as SD1.5 ships it): vocab 49408, 77 positions, width 768,
12 pre-LayerNorm layers of 12 heads (d = 64) with a causal mask, MLP 3072 with quick_gelu, final LayerNorm; the
pipeline consumes ``last_hidden_state`` [B, 77, 768].
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn as nn


@dataclass
class ClipTextConfig:
    vocab_size: int = 49408
    max_position_embeddings: int = 77
    hidden_size: int = 768
    intermediate_size: int = 3072
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    layer_norm_eps: float = 1e-5


class _Embeddings(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.token_embedding = nn.Embedding(c.vocab_size, c.hidden_size)
        self.position_embedding = nn.Embedding(c.max_position_embeddings, c.hidden_size)


class _SelfAttn(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        d = c.hidden_size
        self.heads = c.num_attention_heads
        self.k_proj, self.v_proj, self.q_proj, self.out_proj = nn.Linear(d, d), nn.Linear(d, d), nn.Linear(d, d), nn.Linear(d, d)

    def forward(self, x):
        b, l, c = x.shape
        hd = c // self.heads
        q = self.q_proj(x).view(b, l, self.heads, hd).transpose(1, 2)
        k = self.k_proj(x).view(b, l, self.heads, hd).transpose(1, 2)
        v = self.v_proj(x).view(b, l, self.heads, hd).transpose(1, 2)
        s = q @ k.transpose(-1, -2) / math.sqrt(hd)
        mask = torch.full((l, l), float("-inf")).triu(1)            # causal: token i sees tokens <= i
        p = (s + mask).softmax(-1)
        return self.out_proj((p @ v).transpose(1, 2).reshape(b, l, c))


class _Mlp(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.fc1, self.fc2 = nn.Linear(c.hidden_size, c.intermediate_size), nn.Linear(c.intermediate_size, c.hidden_size)

    def forward(self, x):
        h = self.fc1(x)
        return self.fc2(h * torch.sigmoid(1.702 * h))              # quick_gelu


class _Layer(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.self_attn = _SelfAttn(c)
        self.layer_norm1 = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.mlp = _Mlp(c)
        self.layer_norm2 = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)

    def forward(self, x):
        x = x + self.self_attn(self.layer_norm1(x))
        return x + self.mlp(self.layer_norm2(x))


class _Encoder(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.layers = nn.ModuleList([_Layer(c) for _ in range(c.num_hidden_layers)])


class _TextTransformer(nn.Module):
    def __init__(self, c: ClipTextConfig):
        super().__init__()
        self.embeddings = _Embeddings(c)
        self.encoder = _Encoder(c)
        self.final_layer_norm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)


class OracleClipText(nn.Module):
    def __init__(self, cfg: ClipTextConfig = None):
        super().__init__()
        self.cfg = cfg or ClipTextConfig()
        self.text_model = _TextTransformer(self.cfg)

    @torch.no_grad()
    def forward(self, input_ids: torch.Tensor) -> torch.Tensor:
        """input_ids int64 [B, L<=77] -> last_hidden_state [B, L, 768]."""
        t = self.text_model
        l = input_ids.shape[1]
        x = t.embeddings.token_embedding(input_ids) + t.embeddings.position_embedding(torch.arange(l))[None]
        for layer in t.encoder.layers:
            x = layer(x)
        return t.final_layer_norm(x)


def make_oracle_clip(seed: int = 2, cfg: ClipTextConfig = None, round_to=torch.bfloat16) -> OracleClipText:
    state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    m = OracleClipText(cfg).eval()
    torch.random.set_rng_state(state)
    with torch.no_grad():
        for p in m.parameters():
            if round_to is not None:
                p.copy_(p.to(round_to).float())
            p.requires_grad_(False)
    return m

"""CLIP ViT-L/14 text encoder on the b200ad kernels -- row f2 of SURVEY.md section 8: the step right before the denoise
loop (``encode_prompt`` of AnimateDiffPipeline; the reference loads ``text_encoder`` + ``tokenizer`` inside
``from_pretrained`` at /root/reference/experiments/05_grid_search_ablation.py:130-134 and passes prompt strings at :158-167).

Module tree / state-dict keys follow ``transformers.CLIPTextModel`` (``text_model.embeddings.token_embedding.weight``,
``text_model.encoder.layers.N.self_attn.{q,k,v,out}_proj``, ``...layer_norm{1,2}``, ``...mlp.fc{1,2}``,
``text_model.final_layer_norm``), so a local SD1.5 ``text_encoder/model.safetensors`` loads unchanged.  Per layer:
LayerNorm (b200ad_layernorm) -> fused Q|K|V GEMM (+bias) -> causal 77-token attention (b200ad_attention_small) ->
out-projection GEMM (+bias, +fp32 residual) -> LayerNorm -> fc1 GEMM (+bias) -> quick_gelu (b200ad_eltwise) -> fc2 GEMM
(+bias, +residual); token + position embeddings by b200ad_embed_tokens; final LayerNorm in fp32.  ~40 GFLOP for the two
prompts of a video: off the per-step path, once per prompt.  No CPU fallback.
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from .leaf import LayerNorm, Linear
from .ops import Ops, get_ops
from .unet import _Packable, _dev_f, _dev_w, _mode_of


class _Embeddings(nn.Module):
    def __init__(self, vocab, positions, dim):
        super().__init__()
        self.token_embedding = nn.Embedding(vocab, dim)
        self.position_embedding = nn.Embedding(positions, dim)


class _SelfAttn(nn.Module):
    def __init__(self, dim, heads):
        super().__init__()
        self.heads = heads
        self.k_proj, self.v_proj, self.q_proj, self.out_proj = Linear(dim, dim), Linear(dim, dim), Linear(dim, dim), Linear(dim, dim)


class _Mlp(nn.Module):
    def __init__(self, dim, inter):
        super().__init__()
        self.fc1, self.fc2 = Linear(dim, inter), Linear(inter, dim)


class _Layer(_Packable):
    def __init__(self, dim, heads, inter, eps):
        super().__init__()
        self.self_attn = _SelfAttn(dim, heads)
        self.layer_norm1 = LayerNorm(dim, eps=eps)
        self.mlp = _Mlp(dim, inter)
        self.layer_norm2 = LayerNorm(dim, eps=eps)

    def _pack(self, o: Ops):
        a, m = self.self_attn, self.mlp
        return {"g1": _dev_f(self.layer_norm1.weight), "b1": _dev_f(self.layer_norm1.bias),
                "g2": _dev_f(self.layer_norm2.weight), "b2": _dev_f(self.layer_norm2.bias),
                "wqkv": _dev_w(o, torch.cat([a.q_proj.weight.detach(), a.k_proj.weight.detach(), a.v_proj.weight.detach()], 0)),
                "bqkv": _dev_f(torch.cat([a.q_proj.bias.detach(), a.k_proj.bias.detach(), a.v_proj.bias.detach()], 0)),
                "wo": _dev_w(o, a.out_proj.weight), "bo": _dev_f(a.out_proj.bias),
                "w1": _dev_w(o, m.fc1.weight), "bb1": _dev_f(m.fc1.bias),
                "w2": _dev_w(o, m.fc2.weight), "bb2": _dev_f(m.fc2.bias)}

    def run(self, o: Ops, h, B: int, L: int):
        """h: fp32 [B*L, dim] residual stream, updated in place."""
        p = self.packed(o)
        dim, heads = h.shape[1], self.self_attn.heads
        n = o.layer_norm(h, p["g1"], p["b1"], self.layer_norm1.eps)
        _, qkv = o.gemm(n, p["wqkv"], bias=p["bqkv"], want_op=True)
        att = o.attention_small(qkv[:, :dim], qkv[:, dim:2 * dim], qkv[:, 2 * dim:], B, heads, L, dim // heads, causal=True)
        o.gemm(att, p["wo"], bias=p["bo"], residual=h, out_f32=h)
        n = o.layer_norm(h, p["g2"], p["b2"], self.layer_norm2.eps)
        f1, _ = o.gemm(n, p["w1"], bias=p["bb1"], want_f32=True)
        act = o.quick_gelu(f1)
        o.gemm(o.cast(act, 1, 1, act.shape[0], act.shape[1]), p["w2"], bias=p["bb2"], residual=h, out_f32=h)
        return h


class _Encoder(nn.Module):
    def __init__(self, n, dim, heads, inter, eps):
        super().__init__()
        self.layers = nn.ModuleList([_Layer(dim, heads, inter, eps) for _ in range(n)])


class _TextTransformer(nn.Module):
    def __init__(self, vocab, positions, dim, heads, inter, n, eps):
        super().__init__()
        self.embeddings = _Embeddings(vocab, positions, dim)
        self.encoder = _Encoder(n, dim, heads, inter, eps)
        self.final_layer_norm = LayerNorm(dim, eps=eps)


class CLIPTextEncoder(_Packable):
    """``encoder(input_ids)`` -> last_hidden_state fp32 [B, L, 768] (what AnimateDiffPipeline.encode_prompt consumes)."""

    def __init__(self, vocab_size=49408, max_position_embeddings=77, hidden_size=768, intermediate_size=3072,
                 num_hidden_layers=12, num_attention_heads=12, layer_norm_eps=1e-5, mode: str = "bf16"):
        super().__init__()
        self.text_model = _TextTransformer(vocab_size, max_position_embeddings, hidden_size, num_attention_heads,
                                           intermediate_size, num_hidden_layers, layer_norm_eps)
        self.max_length = max_position_embeddings
        self.eval()
        self.set_mode(mode)
        self.register_load_state_dict_post_hook(lambda m, keys: m.invalidate_all_packed())

    def set_mode(self, mode: str):
        if mode not in ("bf16", "fp32"):
            raise ValueError(mode)
        for m in self.modules():
            m._b200_mode = mode
        return self

    @property
    def mode(self):
        return _mode_of(self)

    def invalidate_all_packed(self):
        for m in self.modules():
            if isinstance(m, _Packable):
                m.invalidate_packed()

    def _apply(self, fn, *a, **kw):
        r = super()._apply(fn, *a, **kw)
        self.invalidate_all_packed()
        return r

    def _pack(self, o: Ops):
        t = self.text_model
        return {"tok": _dev_f(t.embeddings.token_embedding.weight), "pos": _dev_f(t.embeddings.position_embedding.weight),
                "g": _dev_f(t.final_layer_norm.weight), "b": _dev_f(t.final_layer_norm.bias)}

    @torch.no_grad()
    def forward(self, input_ids: torch.Tensor, **kw) -> torch.Tensor:
        if not input_ids.is_cuda:
            raise RuntimeError("CLIPTextEncoder: CUDA tensors required (the b200ad path has no CPU fallback)")
        o = get_ops(self.mode)
        P = self.packed(o)
        B, L = input_ids.shape
        if L > self.max_length:
            raise ValueError(f"sequence length {L} exceeds max_position_embeddings={self.max_length}")
        h = o.embed_tokens(input_ids.long().contiguous(), P["tok"], P["pos"])
        for layer in self.text_model.encoder.layers:
            layer.run(o, h, B, L)
        out = get_ops("fp32").layer_norm(h, P["g"], P["b"], self.text_model.final_layer_norm.eps)
        return out.view(B, L, -1)

    @classmethod
    def from_local(cls, directory: str, mode: str = "bf16"):
        """Load ``<directory>/model.safetensors`` (or ``pytorch_model.bin``) of a local SD1.5 ``text_encoder/`` folder."""
        import json
        cfg_path = os.path.join(directory, "config.json")
        kw = {}
        if os.path.isfile(cfg_path):
            with open(cfg_path) as f:
                c = json.load(f)
            if c.get("hidden_act", "quick_gelu") != "quick_gelu":
                raise NotImplementedError(f"CLIP text encoder with hidden_act={c['hidden_act']!r}")
            kw = {k: c[k] for k in ("vocab_size", "max_position_embeddings", "hidden_size", "intermediate_size",
                                    "num_hidden_layers", "num_attention_heads", "layer_norm_eps") if k in c}
        enc = cls(mode=mode, **kw)
        for name in ("model.safetensors", "model.fp16.safetensors", "pytorch_model.bin"):
            f = os.path.join(directory, name)
            if os.path.isfile(f):
                if name.endswith(".safetensors"):
                    from safetensors.torch import load_file
                    sd = load_file(f)
                else:
                    sd = torch.load(f, map_location="cpu", weights_only=True)
                sd = {k: v for k, v in sd.items() if not k.endswith("position_ids")}
                res = enc.load_state_dict(sd, strict=False)
                if res.missing_keys or res.unexpected_keys:
                    raise RuntimeError(f"text encoder checkpoint does not match: missing {res.missing_keys[:3]}, "
                                       f"unexpected {res.unexpected_keys[:3]}")
                return enc.cuda() if torch.cuda.is_available() else enc
        raise FileNotFoundError(f"no text encoder weights in {directory}")

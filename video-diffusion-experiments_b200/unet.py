"""UNetMotionModel (SD1.5 UNet + AnimateDiff motion adapter v1.5-2) on the b200ad sm_100a kernels.

Surface kept from diffusers as the reference exercises it:
  * ``unet(sample[B,4,F,h,w], timestep, encoder_hidden_states=[B|B*F,77,768]).sample``
    (/root/reference/experiments/03_trace_forward_pass.py:109-115; called per step by AnimateDiffPipeline,
    /root/reference/experiments/05_grid_search_ablation.py:158-167),
  * an ``nn.Module`` tree with diffusers names / state-dict keys (``down_blocks.i.motion_modules.j...``,
    ``...attentions.j.transformer_blocks.0.attn1.to_q``) that owns the fp32 parameters
    (/root/reference/experiments/02_architecture_inspection.py:38-60,98-107),
  * ``attn_processors`` / ``set_attn_processor``.

Execution is B200-first, not a module-by-module translation: activations are channels-last row matrices
``[B*F*h*w, C]`` for the whole network (a 1x1 conv, a Linear over tokens and a Linear over the frame axis are
the same GEMM on the same rows -- no permute is ever materialised), the residual stream is fp32, every tensor
that feeds a contraction is written once in the operand type by the producing kernel's epilogue, and
bias / timestep-embedding / residual / GEGLU are fused into GEMM and conv epilogues.  Every computation below
is a kernel of libb200ad.so (ops.py); torch is used for allocation and views only.  No CPU fallback.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from . import leaf as _leaf
from .leaf import Conv2d, GroupNorm, LayerNorm, Linear, SiLU
from .ops import Ops, _empty, get_ops


@dataclass
class UNetMotionConfig:
    """SD1.5 UNet config + motion adapter v1-5-2 config."""
    in_channels: int = 4
    out_channels: int = 4
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    layers_per_block: int = 2
    num_attention_heads: int = 8
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


@dataclass
class UNetMotionOutput:
    sample: torch.Tensor


@dataclass
class Transformer2DModelOutput:
    sample: torch.Tensor


def _dpad(d: int, o: Ops) -> int:
    """Head stride of q/k/v in the rows the tcgen05 attention kernels read; fp32 mode needs none.  The kernels load
    64-column slabs but only use 16*ceil(d/16) columns of Q/K and d (+1: the ones column) columns of V/O, so the stride
    is just wide enough for those -- 48 / 88 / 168 for d = 40 / 80 / 160 instead of 64 / 128 / 192: a quarter less QKV
    GEMM work and q/k/v traffic.  B200AD_TIGHT_HEADS=0: pad to whole slabs."""
    if o.mode == "fp32":
        return d
    if d > 192:
        raise ValueError(f"head dim {d} > 192 unsupported")
    slab = next(p for p in (64, 128, 192) if d <= p)
    if os.environ.get("B200AD_TIGHT_HEADS", "1") == "0":
        return slab
    return min(slab, max((d + 15) // 16 * 16, (d + 1 + 7) // 8 * 8))


def _dev_w(o: Ops, w: torch.Tensor) -> torch.Tensor:
    return w.detach().to(device="cuda", dtype=o.op_dtype).contiguous()


def _dev_f(v: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    return None if v is None else v.detach().to(device="cuda", dtype=torch.float32).contiguous()


def _pad_heads(w: torch.Tensor, heads: int, dpad: int) -> torch.Tensor:
    """[heads*d, K] -> [heads*dpad, K]; pad rows are zero so the padded q/k/v columns come out as zeros."""
    n, k = w.shape
    d = n // heads
    if d == dpad:
        return w
    out = w.new_zeros(heads, dpad, k)
    out[:, :d] = w.view(heads, d, k)
    return out.view(heads * dpad, k)


class _State:
    """Per-forward context handed down the tree."""
    __slots__ = ("o", "B", "F", "temb", "ctx", "ctx_rows", "kv_div", "kv_cache")

    def __init__(self, o: Ops, B: int, F: int):
        self.o, self.B, self.F = o, B, F
        self.temb = None        # [B, sum(cout)] fp32: all time_emb_proj outputs of the step (one launch)
        self.ctx = None         # operand-typed text rows [Bt*77, 768]
        self.ctx_rows = 0
        self.kv_div = 1         # frames sharing one text row
        self.kv_cache = None    # dict: id(module) -> (kv buffer) for cross-attention K/V of the text


class _Packable(nn.Module):
    """Mixin: operand-typed / re-laid-out copies of the parameters, built on first use per mode."""

    def packed(self, o: Ops):
        cache = self.__dict__.setdefault("_b200_packed", {})
        if o.mode not in cache:
            with torch.no_grad():
                cache[o.mode] = self._pack(o)
        return cache[o.mode]

    def invalidate_packed(self):
        self.__dict__["_b200_packed"] = {}


# ----------------------------------------------------------------------------- attention pieces
class Attention(_Packable):
    """diffusers ``Attention`` (A.2.4): to_q/to_k/to_v without bias, to_out = [Linear(+bias), Dropout]."""

    def __init__(self, query_dim: int, heads: int, cross_attention_dim: Optional[int] = None):
        super().__init__()
        self.heads = heads
        self.inner_dim = query_dim
        self.query_dim = query_dim
        self.cross_attention_dim = cross_attention_dim
        self.is_cross_attention = cross_attention_dim is not None
        kv_dim = query_dim if cross_attention_dim is None else cross_attention_dim
        self.to_q = Linear(query_dim, query_dim, bias=False)
        self.to_k = Linear(kv_dim, query_dim, bias=False)
        self.to_v = Linear(kv_dim, query_dim, bias=False)
        self.to_out = nn.ModuleList([Linear(query_dim, query_dim, bias=True), nn.Dropout(0.0)])
        self.processor = None       # set by UNetMotionModel.set_attn_processor

    def _pack(self, o: Ops):
        return {"wo": _dev_w(o, self.to_out[0].weight), "bo": _dev_f(self.to_out[0].bias)}

    def packed_spatial_self(self, o: Ops):
        p = self.packed(o)
        if "wqkv" not in p:
            dp = _dpad(self.query_dim // self.heads, o)
            w = torch.cat([_pad_heads(self.to_q.weight.detach(), self.heads, dp),
                           _pad_heads(self.to_k.weight.detach(), self.heads, dp),
                           _pad_heads(self.to_v.weight.detach(), self.heads, dp)], 0)
            p["wqkv"], p["dp"] = _dev_w(o, w), dp
            # V pad column d of every head = 1.0 (zero weight row + bias 1): the long self-attention kernel reads
            # its softmax denominator out of the PV product (B200AD_ATTN_V_ONES)
            d = self.query_dim // self.heads
            p["bqkv"] = None
            if dp > d:
                b = torch.zeros(3, self.heads, dp)
                b[2, :, d] = 1.0
                p["bqkv"] = _dev_f(b.reshape(-1))
        return p

    def packed_cross(self, o: Ops):
        p = self.packed(o)
        if "wq" not in p:
            dp = _dpad(self.query_dim // self.heads, o)
            p["wq"] = _dev_w(o, _pad_heads(self.to_q.weight.detach(), self.heads, dp))
            p["wkv"] = _dev_w(o, torch.cat([_pad_heads(self.to_k.weight.detach(), self.heads, dp),
                                            _pad_heads(self.to_v.weight.detach(), self.heads, dp)], 0))
            p["dp"] = dp
        return p

    def packed_temporal(self, o: Ops):
        p = self.packed(o)
        if "wqkv_t" not in p:
            p["wqkv_t"] = _dev_w(o, torch.cat([self.to_q.weight.detach(), self.to_k.weight.detach(),
                                               self.to_v.weight.detach()], 0))
        return p

    # -- the three fused attention bodies; `n_op` = normed input rows (operand type), `h` = fp32 residual stream
    def run_spatial_self(self, st: _State, n_op, h, nimg: int, L: int):
        o = st.o
        p = self.packed_spatial_self(o)
        H, C, dp = self.heads, self.query_dim, p["dp"]
        _, qkv = o.gemm(n_op, p["wqkv"], bias=p["bqkv"], want_op=True)
        hd = H * dp
        att = o.attention(qkv[:, :hd], qkv[:, hd:2 * hd], qkv[:, 2 * hd:], nimg, H, L, L, C // H, dp, 1,
                          v_ones=p["bqkv"] is not None)
        o.gemm(att, p["wo"], bias=p["bo"], residual=h, out_f32=h)
        return h

    def run_cross(self, st: _State, n_op, h, nimg: int, L: int, h_op=None):
        """h_op (optional, operand-typed [rows, C] view): also receives the updated stream (see _fold_tail)."""
        o = st.o
        p = self.packed_cross(o)
        H, C, dp = self.heads, self.query_dim, p["dp"]
        hd = H * dp
        kv = st.kv_cache.get(id(self)) if st.kv_cache is not None else None
        if kv is None:
            _, kv = o.gemm(st.ctx, p["wkv"], want_op=True)          # text K|V: once per forward (or per video)
            if st.kv_cache is not None:
                st.kv_cache[id(self)] = kv
        _, q = o.gemm(n_op, p["wq"], want_op=True)
        Lk = st.ctx_rows // (nimg // st.kv_div)
        att = o.attention(q, kv[:, :hd], kv[:, hd:], nimg, H, L, Lk, C // H, dp, st.kv_div)
        o.gemm(att, p["wo"], bias=p["bo"], residual=h, out_f32=h, out_op=h_op)
        return h

    def run_temporal(self, st: _State, n_op, h, hw: int, h_op=None):
        o = st.o
        p = self.packed_temporal(o)
        _, qkv = o.gemm(n_op, p["wqkv_t"], want_op=True)
        att = o.temporal_attention(qkv, st.B, st.F, hw, self.query_dim, self.heads)
        o.gemm(att, p["wo"], bias=p["bo"], residual=h, out_f32=h, out_op=h_op)
        return h

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **kw):
        """Stand-alone diffusers-style call [Bt, L, C] -> [Bt, L, C] (module mode / attention processors)."""
        from .attention_processor import B200AttnProcessor
        proc = self.processor or B200AttnProcessor()
        return proc(self, hidden_states, encoder_hidden_states=encoder_hidden_states, attention_mask=attention_mask)


class GEGLU(nn.Module):
    def __init__(self, dim_in: int, dim_out: int):
        super().__init__()
        self.proj = Linear(dim_in, dim_out * 2)

    def forward(self, hidden_states, *a, **kw):
        """Module mode: h, g = proj(x).chunk(2, -1); h * gelu(g) (exact erf GELU)."""
        hg = self.proj(hidden_states)
        o = get_ops(_mode_of(self))
        out = o.geglu_gate(_leaf.rows_f32(hg))
        return out.view(*hg.shape[:-1], hg.shape[-1] // 2).to(hidden_states.dtype)


class FeedForward(_Packable):
    """GEGLU feed-forward: net = [GEGLU(C -> 4C), Dropout, Linear(4C -> C)]."""

    def __init__(self, dim: int):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * 4), nn.Dropout(0.0), Linear(dim * 4, dim)])

    def _pack(self, o: Ops):
        w, b = self.net[0].proj.weight.detach(), self.net[0].proj.bias.detach()
        half = w.shape[0] // 2
        # interleave (h_i, g_i) so one 32-column epilogue chunk holds both halves of 16 outputs
        wi = torch.stack([w[:half], w[half:]], 1).reshape(2 * half, -1)
        bi = torch.stack([b[:half], b[half:]], 1).reshape(2 * half)
        return {"w1": _dev_w(o, wi), "b1": _dev_f(bi), "w2": _dev_w(o, self.net[2].weight),
                "b2": _dev_f(self.net[2].bias)}

    def forward(self, hidden_states, *a, **kw):
        """Module mode (diffusers FeedForward.forward): GEGLU -> Dropout -> Linear."""
        for m in self.net:
            hidden_states = m(hidden_states)
        return hidden_states

    def run(self, st: _State, n_op, h, *, op_only: bool, gate_out=None):
        """h + ff(n): returns fp32 (in place) or, when the caller only needs an operand, the operand-typed sum.
        gate_out (operand-typed [rows, 4C] view): only the GEGLU half runs, written there; the second Linear is
        folded into the parent's proj_out GEMM (_fold_tail)."""
        o, p = st.o, self.packed(st.o)
        if gate_out is not None:
            o.gemm(n_op, p["w1"], bias=p["b1"], out_op=gate_out, geglu=True)
            return None
        _, g = o.gemm(n_op, p["w1"], bias=p["b1"], want_op=True, geglu=True)
        if op_only:
            return o.gemm(g, p["w2"], bias=p["b2"], residual=h, want_op=True)[1]
        o.gemm(g, p["w2"], bias=p["b2"], residual=h, out_f32=h)
        return h


class SinusoidalPositionalEmbedding(nn.Module):
    """pe[0,:,0::2] = sin(pos*div), pe[0,:,1::2] = cos(pos*div) (A.2.3); persistent buffer `pe` [1, max_seq, C]."""

    def __init__(self, embed_dim: int, max_seq_length: int = 32):
        super().__init__()
        import math
        position = torch.arange(max_seq_length).unsqueeze(1)
        div_term = torch.exp(torch.arange(0, embed_dim, 2) * (-math.log(10000.0) / embed_dim))
        pe = torch.zeros(1, max_seq_length, embed_dim)
        pe[0, :, 0::2] = torch.sin(position * div_term)
        pe[0, :, 1::2] = torch.cos(position * div_term)
        self.register_buffer("pe", pe)

    def forward(self, x):
        """Module mode: x [Bt, F, C] + pe[:, :F]."""
        _, f, c = x.shape
        return _leaf.add(get_ops(_mode_of(self)), x, self.pe[0, :f].to(x.device), div=1, mod=f * c)


class BasicTransformerBlock(_Packable):
    """Spatial block (self-attn, text cross-attn, GEGLU FF) or temporal block (two self-attentions with
    sinusoidal PE over the frame axis, GEGLU FF)."""

    def __init__(self, dim: int, heads: int, cross_attention_dim: Optional[int], temporal: bool, max_seq: int = 32):
        super().__init__()
        self.temporal = temporal
        self.pos_embed = SinusoidalPositionalEmbedding(dim, max_seq) if temporal else None
        self.norm1 = LayerNorm(dim, eps=1e-5)
        self.attn1 = Attention(dim, heads)
        self.norm2 = LayerNorm(dim, eps=1e-5)
        self.attn2 = Attention(dim, heads, cross_attention_dim=None if temporal else cross_attention_dim)
        self.norm3 = LayerNorm(dim, eps=1e-5)
        self.ff = FeedForward(dim)

    def _pack(self, o: Ops):
        p = {}
        for i, n in enumerate((self.norm1, self.norm2, self.norm3), 1):
            p[f"g{i}"], p[f"b{i}"] = _dev_f(n.weight), _dev_f(n.bias)
        p["pe"] = _dev_f(self.pos_embed.pe[0]) if self.pos_embed is not None else None
        return p

    def forward(self, hidden_states, attention_mask=None, encoder_hidden_states=None, encoder_attention_mask=None,
                timestep=None, cross_attention_kwargs=None, class_labels=None, **kw):
        """Module mode (diffusers BasicTransformerBlock.forward): every sub-module is called, so hooks fire."""
        o = get_ops(_mode_of(self))
        x = hidden_states
        n = self.norm1(x)
        if self.pos_embed is not None:
            n = self.pos_embed(n)
        x = _leaf.add(o, x, self.attn1(n, encoder_hidden_states=None, attention_mask=attention_mask))
        n = self.norm2(x)
        if self.pos_embed is not None:
            n = self.pos_embed(n)
        x = _leaf.add(o, x, self.attn2(n, encoder_hidden_states=None if self.temporal else encoder_hidden_states,
                                       attention_mask=encoder_attention_mask))
        return _leaf.add(o, x, self.ff(self.norm3(x)))

    def run(self, st: _State, h, nimg: int, hw: int, last: bool = True, fold: bool = False):
        """h: fp32 [nimg*hw, C] residual stream (updated in place).  last=True: returns the operand-typed block
        output (its only consumer is the proj_out GEMM) and leaves h WITHOUT the feed-forward term; last=False
        (another block follows): h += ff and h is returned.  last and fold: returns the operand matrix
        [rows, 5C] = (h after attn2 | GEGLU gate) for the parent's folded proj_out GEMM (_fold_tail)."""
        o, p = st.o, self.packed(st.o)
        C = h.shape[1]
        cat = _empty((h.shape[0], 5 * C), h.device, o.op_dtype) if (last and fold) else None
        h_op = None if cat is None else cat[:, :C]
        if self.temporal:
            n = o.layer_norm(h, p["g1"], p["b1"], self.norm1.eps, pe=p["pe"], hw=hw, frames=st.F)
            self.attn1.run_temporal(st, n, h, hw)
            n = o.layer_norm(h, p["g2"], p["b2"], self.norm2.eps, pe=p["pe"], hw=hw, frames=st.F)
            self.attn2.run_temporal(st, n, h, hw, h_op=h_op)
        else:
            n = o.layer_norm(h, p["g1"], p["b1"], self.norm1.eps)
            self.attn1.run_spatial_self(st, n, h, nimg, hw)
            n = o.layer_norm(h, p["g2"], p["b2"], self.norm2.eps)
            self.attn2.run_cross(st, n, h, nimg, hw, h_op=h_op)
        n = o.layer_norm(h, p["g3"], p["b3"], self.norm3.eps)
        if cat is not None:
            self.ff.run(st, n, h, op_only=True, gate_out=cat[:, C:])
            return cat
        return self.ff.run(st, n, h, op_only=last)


def _fold_tail(o: Ops, ff: "FeedForward", w_out: torch.Tensor, b_out: torch.Tensor):
    """The last block's second feed-forward Linear and the transformer's proj_out are two linear maps with nothing
    but the residual add between them:
        x + Wout (h + W2 g + b2) + bout  =  x + [h | g] [Wout | Wout W2]^T + (Wout b2 + bout)
    so they run as ONE GEMM with K = 5C over the operand matrix (h | g) that attn2's out-projection (operand copy of
    the stream) and the GEGLU GEMM write side by side.  Saves the 4 B/element re-read of the fp32 stream, the
    2 B/element write + read of the block output and one launch per transformer.  The composite is formed in fp64 and
    rounded to the operand type once.  bf16 kernel mode only (the fp32 validation mode keeps the module's algebra);
    B200AD_FOLD_TAIL=0 turns it off."""
    if o.mode != "bf16" or os.environ.get("B200AD_FOLD_TAIL", "1") == "0":
        return None, None
    wcat, bcat = fold_tail_weights(ff.net[2].weight, ff.net[2].bias, w_out, b_out)
    return _dev_w(o, wcat), _dev_f(bcat)


def fold_tail_weights(w2, b2, w_out, b_out):
    """fp64 ([Wout | Wout W2] of shape [C, 5C], Wout b2 + bout) -- the algebra of _fold_tail (host side, any device)."""
    w2, b2 = w2.detach().double(), b2.detach().double()
    wo, bo = w_out.detach().double().reshape(w_out.shape[0], -1), b_out.detach().double()
    return torch.cat([wo, wo @ w2], 1), wo @ b2 + bo


class Transformer2DModel(_Packable):
    """Spatial transformer (A.2.2): GN32(eps 1e-6) -> 1x1 conv -> block -> 1x1 conv -> + residual."""

    def __init__(self, channels: int, heads: int, cross_attention_dim: int, groups: int = 32):
        super().__init__()
        self.norm = GroupNorm(groups, channels, eps=1e-6)
        self.proj_in = Conv2d(channels, channels, 1)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(channels, heads, cross_attention_dim, temporal=False)])
        self.proj_out = Conv2d(channels, channels, 1)

    def _pack(self, o: Ops):
        c = self.proj_in.weight.shape[0]
        wcat, bcat = _fold_tail(o, self.transformer_blocks[-1].ff, self.proj_out.weight, self.proj_out.bias)
        return {"gn_g": _dev_f(self.norm.weight), "gn_b": _dev_f(self.norm.bias),
                "win": _dev_w(o, self.proj_in.weight.reshape(c, c)), "bin": _dev_f(self.proj_in.bias),
                "wout": _dev_w(o, self.proj_out.weight.reshape(c, c)), "bout": _dev_f(self.proj_out.bias),
                "wcat": wcat, "bcat": bcat}

    def forward(self, hidden_states, encoder_hidden_states=None, timestep=None, added_cond_kwargs=None,
                class_labels=None, cross_attention_kwargs=None, attention_mask=None, encoder_attention_mask=None,
                return_dict: bool = True):
        """Module mode (diffusers Transformer2DModel.forward, use_linear_projection=False)."""
        n, c, h, w = hidden_states.shape
        residual = hidden_states
        x = self.proj_in(self.norm(hidden_states))
        x = x.permute(0, 2, 3, 1).reshape(n, h * w, c)
        for blk in self.transformer_blocks:
            x = blk(x, attention_mask=attention_mask, encoder_hidden_states=encoder_hidden_states,
                    encoder_attention_mask=encoder_attention_mask, timestep=timestep)
        x = x.reshape(n, h, w, c).permute(0, 3, 1, 2).contiguous()
        out = _leaf.add(get_ops(_mode_of(self)), self.proj_out(x), residual)
        return Transformer2DModelOutput(sample=out) if return_dict else (out,)

    def run(self, st: _State, x, nimg: int, hw: int):
        """x: fp32 [nimg*hw, C], updated in place (x += proj_out(block(proj_in(GN(x)))))."""
        o, p = st.o, self.packed(st.o)
        a = o.group_norm(x, None, nimg, hw, self.norm.num_groups, p["gn_g"], p["gn_b"], self.norm.eps, silu=False)
        h, _ = o.gemm(a, p["win"], bias=p["bin"], want_f32=True)
        for blk in self.transformer_blocks[:-1]:
            blk.run(st, h, nimg, hw, last=False)
        if p["wcat"] is not None:
            cat = self.transformer_blocks[-1].run(st, h, nimg, hw, fold=True)
            o.gemm(cat, p["wcat"], bias=p["bcat"], residual=x, out_f32=x, stats_hw=hw)
            return x
        hb = self.transformer_blocks[-1].run(st, h, nimg, hw)
        o.gemm(hb, p["wout"], bias=p["bout"], residual=x, out_f32=x, stats_hw=hw)     # x feeds a GroupNorm next
        return x


class AnimateDiffTransformer3D(_Packable):
    """Motion module (A.2.3).  GroupNorm statistics span all F frames of a batch item; the temporal
    transformer works on the same (b, f, hw)-ordered rows as the rest of the network: Linear layers are
    row-wise, and the attention / PE kernels index the frame axis in place (no permute copies)."""

    def __init__(self, channels: int, heads: int, groups: int = 32, max_seq: int = 32):
        super().__init__()
        self.norm = GroupNorm(groups, channels, eps=1e-6)
        self.proj_in = Linear(channels, channels)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(channels, heads, None, temporal=True, max_seq=max_seq)])
        self.proj_out = Linear(channels, channels)

    def _pack(self, o: Ops):
        wcat, bcat = _fold_tail(o, self.transformer_blocks[-1].ff, self.proj_out.weight, self.proj_out.bias)
        return {"gn_g": _dev_f(self.norm.weight), "gn_b": _dev_f(self.norm.bias),
                "win": _dev_w(o, self.proj_in.weight), "bin": _dev_f(self.proj_in.bias),
                "wout": _dev_w(o, self.proj_out.weight), "bout": _dev_f(self.proj_out.bias),
                "wcat": wcat, "bcat": bcat}

    def run(self, st: _State, x, hw: int):
        o, p = st.o, self.packed(st.o)
        nimg = st.B * st.F
        a = o.group_norm(x, None, nimg, hw, self.norm.num_groups, p["gn_g"], p["gn_b"], self.norm.eps, silu=False,
                         frames_per_stat=st.F)
        h, _ = o.gemm(a, p["win"], bias=p["bin"], want_f32=True)
        for blk in self.transformer_blocks[:-1]:
            blk.run(st, h, nimg, hw, last=False)
        if p["wcat"] is not None:
            cat = self.transformer_blocks[-1].run(st, h, nimg, hw, fold=True)
            o.gemm(cat, p["wcat"], bias=p["bcat"], residual=x, out_f32=x, stats_hw=hw)
            return x
        hb = self.transformer_blocks[-1].run(st, h, nimg, hw)
        o.gemm(hb, p["wout"], bias=p["bout"], residual=x, out_f32=x, stats_hw=hw)     # x feeds a GroupNorm next
        return x

    def forward(self, hidden_states, encoder_hidden_states=None, timestep=None, class_labels=None,
                num_frames: int = 1, cross_attention_kwargs=None, **kw):
        """Module mode (diffusers AnimateDiffTransformer3D.forward): [B*F, C, h, w] -> same, through the sub-modules with
        diffusers' tensor shapes -- the temporal attention sees [B*h*w, F, C]
        (/root/reference/experiments/03_trace_forward_pass.py:160-169,182-205; docs/02_video_diffusion_architecture.md:69-81)."""
        bf, c, h, w = hidden_states.shape
        b = bf // num_frames
        residual = hidden_states
        x = hidden_states[None, :].reshape(b, num_frames, c, h, w).permute(0, 2, 1, 3, 4)
        x = self.norm(x)
        x = x.permute(0, 3, 4, 2, 1).reshape(b * h * w, num_frames, c)
        x = self.proj_in(x)
        for blk in self.transformer_blocks:
            x = blk(x, encoder_hidden_states=encoder_hidden_states, timestep=timestep)
        x = self.proj_out(x)
        x = x[None, None, :].reshape(b, h, w, num_frames, c).permute(0, 3, 4, 1, 2).reshape(bf, c, h, w).contiguous()
        return _leaf.add(get_ops(_mode_of(self)), x, residual)


def _mode_of(m: nn.Module) -> str:
    return getattr(m, "_b200_mode", "bf16")


def _pack_upconv3x3(o: Ops, w: torch.Tensor) -> torch.Tensor:
    """torch [cout, cin, 3, 3] -> [4, cout, 4*cin] for b200ad_upconv3x3: phase (py, px) of the nearest-x2 upsampled
    image sees a 2x2 convolution of the low-resolution image whose tap (ty, tx) is the sum of the 3x3 weights that land
    on source pixel (y + ty - 1 + py, x + tx - 1 + px): rows ky with floor((py + ky - 1) / 2) == ty - 1 + py.
    Summed in fp64, rounded to the operand type once; k = (ty*2 + tx)*cin + c."""
    return _dev_w(o, upconv_weights(w))


def upconv_weights(w: torch.Tensor) -> torch.Tensor:
    """fp64 [4, cout, 4*cin] sub-pixel weights of a [cout, cin, 3, 3] convolution (host side, any device)."""
    wd = w.detach().double()
    cout, cin = wd.shape[:2]
    rows = {0: ((0,), (1, 2)), 1: ((0, 1), (2,))}      # phase -> (ky of tap 0, ky of tap 1)
    out = torch.zeros(4, cout, 2, 2, cin, dtype=torch.float64, device=wd.device)
    for py in (0, 1):
        for px in (0, 1):
            for ty in (0, 1):
                for tx in (0, 1):
                    for ky in rows[py][ty]:
                        for kx in rows[px][tx]:
                            out[py * 2 + px, :, ty, tx] += wd[:, :, ky, kx]
    return out.reshape(4, cout, 4 * cin)


def _pack_conv3x3(o: Ops, w: torch.Tensor) -> torch.Tensor:
    """torch [cout, cin, 3, 3] -> [cout, 9*cin] with k = (ky*3+kx)*cin + c (one K block per filter tap)."""
    return _dev_w(o, w.detach().permute(0, 2, 3, 1).reshape(w.shape[0], -1))


# ----------------------------------------------------------------------------- conv pieces
class ResnetBlock2D(_Packable):
    """A.2.1: GN32+SiLU -> conv3x3 (+bias +temb) -> GN32+SiLU -> conv3x3 (+bias +shortcut/identity residual)."""

    def __init__(self, cin: int, cout: int, temb_dim: int, groups: int = 32, eps: float = 1e-5):
        super().__init__()
        self.in_channels, self.out_channels = cin, cout
        self.norm1 = GroupNorm(groups, cin, eps=eps)
        self.conv1 = Conv2d(cin, cout, 3, padding=1)
        self.time_emb_proj = Linear(temb_dim, cout)
        self.norm2 = GroupNorm(groups, cout, eps=eps)
        self.dropout = nn.Dropout(0.0)
        self.conv2 = Conv2d(cout, cout, 3, padding=1)
        self.nonlinearity = SiLU()
        self.conv_shortcut = Conv2d(cin, cout, 1) if cin != cout else None
        self._temb_off = None       # column offset into the step's fused time_emb_proj output

    def _pack(self, o: Ops):
        p = {"g1": _dev_f(self.norm1.weight), "b1": _dev_f(self.norm1.bias),
             "g2": _dev_f(self.norm2.weight), "b2": _dev_f(self.norm2.bias),
             "w1": _pack_conv3x3(o, self.conv1.weight), "cb1": _dev_f(self.conv1.bias),
             "w2": _pack_conv3x3(o, self.conv2.weight), "cb2": _dev_f(self.conv2.bias)}
        if self.conv_shortcut is not None:
            w = self.conv_shortcut.weight
            p["wsc"], p["bsc"] = _dev_w(o, w.reshape(w.shape[0], w.shape[1])), _dev_f(self.conv_shortcut.bias)
            # bf16 mode: the 1x1 shortcut rides along as a second K segment of conv2's implicit GEMM
            # (b200ad_conv3x3_sc), its fp32 result is never written / re-read.  B200AD_CONV_SC=0 keeps two launches.
            if o.mode == "bf16" and self.in_channels % 64 == 0 and os.environ.get("B200AD_CONV_SC", "1") != "0":
                p["w2sc"] = torch.cat([p["w2"], p["wsc"]], 1).contiguous()
                p["b2sc"] = p["cb2"] + p["bsc"]
        return p

    def run(self, st: _State, x, skip, nimg: int, h: int, w: int, temb_proj=None):
        """x (and optional skip: the UNet's torch.cat) fp32 rows -> new fp32 rows [nimg*h*w, cout]."""
        o, p = st.o, self.packed(st.o)
        hw = h * w
        G = self.norm1.num_groups
        if temb_proj is None:
            temb_proj = st.temb[:, self._temb_off:self._temb_off + self.out_channels]
        rb_div = (nimg // temb_proj.shape[0]) * hw          # rows sharing one timestep-embedding row
        need_raw = self.conv_shortcut is not None
        a = o.group_norm(x, skip, nimg, hw, G, p["g1"], p["b1"], self.norm1.eps, silu=True, want_raw=need_raw)
        raw = None
        if need_raw:
            a, raw = a
        h32, _ = o.conv3x3(a, p["w1"], nimg, h, w, bias=p["cb1"], row_bias=temb_proj, rb_div=rb_div, stats=True)
        a2 = o.group_norm(h32, None, nimg, hw, G, p["g2"], p["b2"], self.norm2.eps, silu=True)
        if need_raw and "w2sc" in p:
            return o.conv3x3_sc(a2, p["w2sc"], raw, nimg, h, w, bias=p["b2sc"], stats=True)
        if need_raw:
            res, _ = o.gemm(raw, p["wsc"], bias=p["bsc"], want_f32=True)
        else:
            res = x
        out, _ = o.conv3x3(a2, p["w2"], nimg, h, w, bias=p["cb2"], residual=res, stats=True)
        return out

    def forward(self, input_tensor, temb, *a, **kw):
        """Module mode (diffusers ResnetBlock2D.forward): NCHW in/out, temb [N, 1280] (repeated per frame)."""
        o = get_ops(_mode_of(self))
        n, _, hh, ww = input_tensor.shape
        h = self.conv1(self.nonlinearity(self.norm1(input_tensor)))
        t = self.time_emb_proj(self.nonlinearity(temb))                     # [N, cout]
        h = _leaf.add(o, h, t, div=hh * ww, mod=t.numel())                  # h + t[:, :, None, None]
        h = self.conv2(self.dropout(self.nonlinearity(self.norm2(h))))
        x = input_tensor if self.conv_shortcut is None else self.conv_shortcut(input_tensor)
        return _leaf.add(o, x, h)


class Downsample2D(_Packable):
    """Conv2d(C, C, 3, stride 2, pad 1): space-to-depth phase split, then the same implicit-GEMM conv."""

    def __init__(self, channels: int):
        super().__init__()
        self.conv = Conv2d(channels, channels, 3, stride=2, padding=1)

    def _pack(self, o: Ops):
        return {"w": _pack_conv3x3(o, self.conv.weight), "b": _dev_f(self.conv.bias)}

    def forward(self, hidden_states, *a, **kw):
        return self.conv(hidden_states)

    def run(self, st: _State, x, nimg: int, h: int, w: int):
        o, p = st.o, self.packed(st.o)
        s2d = o.cast(x, nimg, h, w, x.shape[1], mode=1)
        out, _ = o.conv3x3(s2d, p["w"], nimg, h // 2, w // 2, stride2=True, bias=p["b"], stats=True)
        return out


class Upsample2D(_Packable):
    """nearest x2 (written once in the operand type) then Conv2d(C, C, 3, pad 1)."""

    def __init__(self, channels: int):
        super().__init__()
        self.conv = Conv2d(channels, channels, 3, padding=1)

    def _pack(self, o: Ops):
        # bf16 mode: the upsampled image is never formed -- four 2x2 sub-pixel convolutions over the low-resolution
        # image (4/9 of the MACs; b200ad_upconv3x3).  B200AD_UPCONV=0 (and the fp32 validation mode) keep the direct form.
        sub = o.mode == "bf16" and os.environ.get("B200AD_UPCONV", "1") != "0"
        return {"w": None if sub else _pack_conv3x3(o, self.conv.weight), "b": _dev_f(self.conv.bias),
                "w4": _pack_upconv3x3(o, self.conv.weight) if sub else None}

    def forward(self, hidden_states, output_size=None, *a, **kw):
        """Module mode: F.interpolate(scale_factor=2, mode="nearest") (b200ad_cast mode 2) then the conv."""
        n, c, h, w = hidden_states.shape
        o = get_ops("fp32")
        up = o.cast(_leaf.nchw_to_rows(hidden_states), n, h, w, c, mode=2)
        return self.conv(_leaf.rows_to_nchw(up, n, 2 * h, 2 * w, hidden_states.dtype))

    def run(self, st: _State, x, nimg: int, h: int, w: int):
        o, p = st.o, self.packed(st.o)
        if p["w4"] is not None:
            return o.upconv3x3(o.cast(x, nimg, h, w, x.shape[1]), p["w4"], nimg, h, w, bias=p["b"], stats=True)
        up = o.cast(x, nimg, h, w, x.shape[1], mode=2)
        out, _ = o.conv3x3(up, p["w"], nimg, 2 * h, 2 * w, bias=p["b"], stats=True)
        return out


# ----------------------------------------------------------------------------- blocks
class DownBlockMotion(nn.Module):
    def __init__(self, cfg: UNetMotionConfig, cin: int, cout: int, cross: bool, add_downsample: bool):
        super().__init__()
        td = cfg.time_embed_dim
        self.resnets = nn.ModuleList()
        self.motion_modules = nn.ModuleList()
        if cross:
            self.attentions = nn.ModuleList()
        self.has_cross_attention = cross
        for i in range(cfg.layers_per_block):
            self.resnets.append(ResnetBlock2D(cin if i == 0 else cout, cout, td, cfg.norm_num_groups, cfg.norm_eps))
            if cross:
                self.attentions.append(Transformer2DModel(cout, cfg.num_attention_heads, cfg.cross_attention_dim,
                                                          cfg.norm_num_groups))
            self.motion_modules.append(AnimateDiffTransformer3D(cout, cfg.motion_num_attention_heads,
                                                                cfg.motion_norm_num_groups, cfg.motion_max_seq_length))
        self.downsamplers = nn.ModuleList([Downsample2D(cout)]) if add_downsample else None

    def forward(self, hidden_states, temb=None, encoder_hidden_states=None, attention_mask=None, num_frames: int = 1,
                encoder_attention_mask=None, cross_attention_kwargs=None, additional_residuals=None, **kw):
        """Module mode (diffusers [CrossAttn]DownBlockMotion.forward) -> (hidden_states, output_states)."""
        x, outs = hidden_states, ()
        for i, res in enumerate(self.resnets):
            x = res(x, temb)
            if self.has_cross_attention:
                x = self.attentions[i](x, encoder_hidden_states=encoder_hidden_states, return_dict=False)[0]
            x = self.motion_modules[i](x, num_frames=num_frames)
            outs += (x,)
        if self.downsamplers is not None:
            for d in self.downsamplers:
                x = d(x)
            outs += (x,)
        return x, outs

    def run(self, st: _State, x, h: int, w: int):
        nimg = st.B * st.F
        skips = []
        for i, res in enumerate(self.resnets):
            x = res.run(st, x, None, nimg, h, w)
            if self.has_cross_attention:
                x = self.attentions[i].run(st, x, nimg, h * w)
            x = self.motion_modules[i].run(st, x, h * w)
            skips.append(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0].run(st, x, nimg, h, w)
            h, w = h // 2, w // 2
            skips.append(x)
        return x, skips, h, w


class CrossAttnDownBlockMotion(DownBlockMotion):
    pass


class UNetMidBlockCrossAttnMotion(nn.Module):
    def __init__(self, cfg: UNetMotionConfig, c: int):
        super().__init__()
        td = cfg.time_embed_dim
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, td, cfg.norm_num_groups, cfg.norm_eps),
                                      ResnetBlock2D(c, c, td, cfg.norm_num_groups, cfg.norm_eps)])
        self.attentions = nn.ModuleList([Transformer2DModel(c, cfg.num_attention_heads, cfg.cross_attention_dim,
                                                            cfg.norm_num_groups)])
        self.motion_modules = nn.ModuleList([AnimateDiffTransformer3D(c, cfg.motion_num_attention_heads,
                                                                      cfg.motion_norm_num_groups,
                                                                      cfg.motion_max_seq_length)])

    def forward(self, hidden_states, temb=None, encoder_hidden_states=None, attention_mask=None,
                cross_attention_kwargs=None, encoder_attention_mask=None, num_frames: int = 1, **kw):
        """Module mode (diffusers UNetMidBlockCrossAttnMotion.forward)."""
        x = self.resnets[0](hidden_states, temb)
        x = self.attentions[0](x, encoder_hidden_states=encoder_hidden_states, return_dict=False)[0]
        x = self.motion_modules[0](x, num_frames=num_frames)
        return self.resnets[1](x, temb)

    def run(self, st: _State, x, h: int, w: int):
        nimg = st.B * st.F
        x = self.resnets[0].run(st, x, None, nimg, h, w)
        x = self.attentions[0].run(st, x, nimg, h * w)
        x = self.motion_modules[0].run(st, x, h * w)
        return self.resnets[1].run(st, x, None, nimg, h, w)


class UpBlockMotion(nn.Module):
    def __init__(self, cfg: UNetMotionConfig, prev_c: int, out_c: int, skip_cs: List[int], cross: bool,
                 add_upsample: bool):
        super().__init__()
        td = cfg.time_embed_dim
        self.resnets = nn.ModuleList()
        self.motion_modules = nn.ModuleList()
        if cross:
            self.attentions = nn.ModuleList()
        self.has_cross_attention = cross
        for i, sc in enumerate(skip_cs):
            cin = (prev_c if i == 0 else out_c) + sc
            self.resnets.append(ResnetBlock2D(cin, out_c, td, cfg.norm_num_groups, cfg.norm_eps))
            if cross:
                self.attentions.append(Transformer2DModel(out_c, cfg.num_attention_heads, cfg.cross_attention_dim,
                                                          cfg.norm_num_groups))
            self.motion_modules.append(AnimateDiffTransformer3D(out_c, cfg.motion_num_attention_heads,
                                                                cfg.motion_norm_num_groups, cfg.motion_max_seq_length))
        self.upsamplers = nn.ModuleList([Upsample2D(out_c)]) if add_upsample else None

    def forward(self, hidden_states, res_hidden_states_tuple, temb=None, encoder_hidden_states=None,
                upsample_size=None, attention_mask=None, encoder_attention_mask=None, cross_attention_kwargs=None,
                num_frames: int = 1, **kw):
        """Module mode (diffusers [CrossAttn]UpBlockMotion.forward): skips are popped from the END of the tuple."""
        x = hidden_states
        for i, res in enumerate(self.resnets):
            skip = res_hidden_states_tuple[-1]
            res_hidden_states_tuple = res_hidden_states_tuple[:-1]
            x = res(torch.cat([x, skip], dim=1), temb)
            if self.has_cross_attention:
                x = self.attentions[i](x, encoder_hidden_states=encoder_hidden_states, return_dict=False)[0]
            x = self.motion_modules[i](x, num_frames=num_frames)
        if self.upsamplers is not None:
            for u in self.upsamplers:
                x = u(x, upsample_size)
        return x

    def run(self, st: _State, x, skips: list, h: int, w: int):
        nimg = st.B * st.F
        for i, res in enumerate(self.resnets):
            x = res.run(st, x, skips.pop(), nimg, h, w)      # torch.cat([x, skip], 1) happens inside GroupNorm
            if self.has_cross_attention:
                x = self.attentions[i].run(st, x, nimg, h * w)
            x = self.motion_modules[i].run(st, x, h * w)
        if self.upsamplers is not None:
            x = self.upsamplers[0].run(st, x, nimg, h, w)
            h, w = 2 * h, 2 * w
        return x, h, w


class CrossAttnUpBlockMotion(UpBlockMotion):
    pass


class Timesteps(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.num_channels = dim

    def forward(self, timesteps):
        """Module mode: sinusoid(dim), flip_sin_to_cos=True, freq_shift=0 -> fp32 [B, dim]."""
        t = timesteps.to(device="cuda", dtype=torch.float32).reshape(-1).contiguous()
        return get_ops("fp32").timestep_sinusoid(t, self.num_channels)


class TimestepEmbedding(nn.Module):
    def __init__(self, cin: int, dim: int):
        super().__init__()
        self.linear_1 = Linear(cin, dim)
        self.act = SiLU()
        self.linear_2 = Linear(dim, dim)

    def forward(self, sample, condition=None):
        return self.linear_2(self.act(self.linear_1(sample)))


# ----------------------------------------------------------------------------- the UNet
class UNetMotionModel(_Packable):
    """forward(sample[B,4,F,h,w], timestep, encoder_hidden_states) -> UNetMotionOutput(sample=[B,4,F,h,w]).

    ``mode``: 'bf16' (tcgen05 tensor cores, default) or 'fp32' (SIMT kernels; the north star's fp32 kernel mode).
    """

    def __init__(self, cfg: Optional[UNetMotionConfig] = None, mode: str = "bf16"):
        super().__init__()
        cfg = cfg or UNetMotionConfig()
        self.cfg = cfg
        self.config = cfg
        boc = cfg.block_out_channels
        self.conv_in = Conv2d(cfg.in_channels, boc[0], 3, padding=1)
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
            scs = [skip_channels.pop() for _ in range(cfg.layers_per_block + 1)]
            cls = CrossAttnUpBlockMotion if rev_cross[i] else UpBlockMotion
            self.up_blocks.append(cls(cfg, prev, oc, scs, rev_cross[i], add_upsample=not last))
            prev = oc
        self.conv_norm_out = GroupNorm(cfg.norm_num_groups, boc[0], eps=cfg.norm_eps)
        self.conv_act = SiLU()
        self.conv_out = Conv2d(boc[0], cfg.out_channels, 3, padding=1)
        self.set_mode(mode)
        self.eval()
        self._kv_cache: Optional[Dict[int, torch.Tensor]] = None
        self._kv_cache_key = None
        self.register_load_state_dict_post_hook(lambda m, keys: m.invalidate_all_packed())

    # ---- configuration -------------------------------------------------------------------------
    def set_mode(self, mode: str):
        if mode not in ("bf16", "fp32"):
            raise ValueError(mode)
        for m in self.modules():
            m._b200_mode = mode
        return self

    @property
    def mode(self) -> str:
        return _mode_of(self)

    @property
    def dtype(self):
        return next(self.parameters()).dtype

    @property
    def device(self):
        return next(self.parameters()).device

    def invalidate_all_packed(self):
        for m in self.modules():
            if isinstance(m, _Packable):
                m.invalidate_packed()
        self._kv_cache, self._kv_cache_key = None, None

    def _apply(self, fn, *a, **kw):          # .to()/.half()/.cuda() change the master copy -> repack lazily
        r = super()._apply(fn, *a, **kw)
        self.invalidate_all_packed()
        return r

    @property
    def attn_processors(self) -> Dict[str, object]:
        from .attention_processor import B200AttnProcessor
        return {f"{n}.processor": (m.processor or B200AttnProcessor())
                for n, m in self.named_modules() if isinstance(m, Attention)}

    def set_attn_processor(self, processor):
        mods = [(n, m) for n, m in self.named_modules() if isinstance(m, Attention)]
        if isinstance(processor, dict):
            if len(processor) != len(mods):
                raise ValueError(f"a dict of {len(processor)} processors was passed for {len(mods)} attention layers")
            for n, m in mods:
                m.processor = processor[f"{n}.processor"]
        else:
            for _, m in mods:
                m.processor = processor

    # ---- packing ---------------------------------------------------------------------------------
    def _resnets(self) -> List[ResnetBlock2D]:
        return [m for m in self.modules() if isinstance(m, ResnetBlock2D)]

    def _pack(self, o: Ops):
        ws, bs, off = [], [], 0
        for r in self._resnets():
            r._temb_off = off
            off += r.out_channels
            ws.append(r.time_emb_proj.weight.detach())
            bs.append(r.time_emb_proj.bias.detach())
        te = self.time_embedding
        return {"temb_w": _dev_f(torch.cat(ws, 0)), "temb_b": _dev_f(torch.cat(bs, 0)),
                "t1w": _dev_f(te.linear_1.weight), "t1b": _dev_f(te.linear_1.bias),
                "t2w": _dev_f(te.linear_2.weight), "t2b": _dev_f(te.linear_2.bias),
                "cin_w": _dev_f(self.conv_in.weight), "cin_b": _dev_f(self.conv_in.bias),
                "cout_w": _dev_f(self.conv_out.weight), "cout_b": _dev_f(self.conv_out.bias),
                "gn_g": _dev_f(self.conv_norm_out.weight), "gn_b": _dev_f(self.conv_norm_out.bias)}

    def prepack(self):
        """Build every packed weight now (otherwise done lazily on the first forward)."""
        o = get_ops(self.mode)
        self.packed(o)
        for m in self.modules():
            if isinstance(m, Attention):
                continue
            if isinstance(m, _Packable) and m is not self:
                m.packed(o)
        for m in self.modules():
            if isinstance(m, BasicTransformerBlock):
                if m.temporal:
                    m.attn1.packed_temporal(o), m.attn2.packed_temporal(o)
                else:
                    m.attn1.packed_spatial_self(o), m.attn2.packed_cross(o)
        return self

    # ---- forward -----------------------------------------------------------------------------------
    def _timestep_tensor(self, timestep, B: int) -> torch.Tensor:
        t = timestep
        if not torch.is_tensor(t):
            t = torch.tensor([float(t)], dtype=torch.float32)
        t = t.to(device="cuda", dtype=torch.float32).reshape(-1)
        if t.numel() == 1 and B > 1:
            t = t.expand(B)
        if t.numel() != B:
            raise ValueError(f"timestep has {t.numel()} entries for batch {B}")
        return t.contiguous()

    def cache_text_kv(self, enable: bool = True):
        """Keep the cross-attention K/V projections of the text between forwards with the same
        encoder_hidden_states tensor (they do not depend on the timestep)."""
        self._kv_cache = {} if enable else None
        self._kv_cache_key = None

    @torch.no_grad()
    def text_kv(self, encoder_hidden_states, out: Optional[dict] = None) -> dict:
        """Cross-attention K|V projections of the text for all 16 spatial transformers (they depend on neither
        the timestep nor the latents: once per video).  With `out` (a previous result) the buffers are
        overwritten in place, so a captured CUDA graph that reads them sees the new prompt."""
        o = get_ops(self.mode)
        e2 = encoder_hidden_states.reshape(-1, encoder_hidden_states.shape[-1]).float().contiguous()
        ctx = o.cast(e2, 1, 1, e2.shape[0], e2.shape[1], mode=0)
        res = out if out is not None else {}
        for m in self.modules():
            if isinstance(m, Attention) and m.is_cross_attention:
                p = m.packed_cross(o)
                res[id(m)] = o.gemm(ctx, p["wkv"], out_op=res.get(id(m)), want_op=True)[1]
        res["rows"] = e2.shape[0]
        return res

    # ---- module mode ------------------------------------------------------------------------------
    module_mode = False          # True: always run through the sub-modules (diffusers' eager structure)

    def _hooks_registered(self) -> bool:
        """True when any SUB-module carries a forward (pre-)hook: the fused plan never calls sub-modules, so a
        tracer (/root/reference/utils/forward_tracer.py:177-194) would see nothing."""
        import torch.nn.modules.module as _m
        if _m._global_forward_hooks or _m._global_forward_pre_hooks:
            return True
        for m in self.modules():
            if m is not self and (m._forward_hooks or m._forward_pre_hooks):
                return True
        return False

    @torch.no_grad()
    def _forward_modules(self, sample, timestep, encoder_hidden_states):
        """diffusers' UNetMotionModel.forward, module by module: every sub-module is invoked
        through ``module(...)`` with diffusers' tensor shapes, each running its own libb200ad kernel (leaf.py)."""
        B, _, F, h, w = sample.shape
        t = self._timestep_tensor(timestep, B)
        emb = self.time_embedding(self.time_proj(t).to(sample.dtype))
        emb = emb.repeat_interleave(F, dim=0)
        ctx = encoder_hidden_states.to(sample.dtype)
        if ctx.shape[0] == B:
            ctx = ctx.repeat_interleave(F, dim=0)
        elif ctx.shape[0] != B * F:
            raise ValueError(f"encoder_hidden_states batch {ctx.shape[0]} is neither B={B} nor B*F={B * F}")
        x = sample.permute(0, 2, 1, 3, 4).reshape(B * F, -1, h, w)
        x = self.conv_in(x)
        skips = (x,)
        for blk in self.down_blocks:
            x, s = blk(x, temb=emb, encoder_hidden_states=ctx, num_frames=F)
            skips += s
        x = self.mid_block(x, temb=emb, encoder_hidden_states=ctx, num_frames=F)
        for blk in self.up_blocks:
            n = len(blk.resnets)
            x = blk(x, skips[-n:], temb=emb, encoder_hidden_states=ctx, num_frames=F)
            skips = skips[:-n]
        x = self.conv_out(self.conv_act(self.conv_norm_out(x)))
        x = x[None, :].reshape(B, F, -1, h, w).permute(0, 2, 1, 3, 4)
        return UNetMotionOutput(sample=x.contiguous())

    @torch.no_grad()
    def forward(self, sample, timestep, encoder_hidden_states, text_kv=None, **kw):
        if not sample.is_cuda:
            raise RuntimeError("UNetMotionModel: CUDA tensors required (the b200ad path has no CPU fallback)")
        if sample.shape[2] > self.cfg.motion_max_seq_length:
            raise ValueError(f"num_frames={sample.shape[2]} exceeds motion_max_seq_length={self.cfg.motion_max_seq_length}")
        if text_kv is None and (self.module_mode or self._hooks_registered()):
            return self._forward_modules(sample, timestep, encoder_hidden_states)
        o = get_ops(self.mode)
        o.colstats.clear()
        P = self.packed(o)
        B, cin, F, h, w = sample.shape
        if F > self.cfg.motion_max_seq_length:
            raise ValueError(f"num_frames={F} exceeds motion_max_seq_length={self.cfg.motion_max_seq_length}")
        st = _State(o, B, F)
        nimg = B * F
        # text rows in the operand type; [B,77,768] (diffusers <= 0.30) or [B*F,77,768] (>= 0.31)
        ehs = encoder_hidden_states
        if ehs.shape[0] == B:
            st.kv_div = F
        elif ehs.shape[0] == nimg:
            st.kv_div = 1
        else:
            raise ValueError(f"encoder_hidden_states batch {ehs.shape[0]} is neither B={B} nor B*F={nimg}")
        if text_kv is not None:                       # caller-owned K/V buffers (DenoisePlan / CUDA graph)
            st.kv_cache, st.ctx_rows = text_kv, text_kv["rows"]
            if st.ctx_rows != ehs.shape[0] * ehs.shape[1]:
                raise ValueError("text_kv was computed for a different encoder_hidden_states shape")
        else:
            if self._kv_cache is not None:
                key = (ehs.data_ptr(), ehs._version, tuple(ehs.shape), self.mode)
                if key != self._kv_cache_key:
                    self._kv_cache, self._kv_cache_key = {}, key
                st.kv_cache = self._kv_cache
            e2 = ehs.reshape(-1, ehs.shape[-1]).float().contiguous()
            st.ctx = o.cast(e2, 1, 1, e2.shape[0], e2.shape[1], mode=0)
            st.ctx_rows = e2.shape[0]
        # time embedding: sinusoid -> Linear -> SiLU -> Linear, then every resnet's Linear(SiLU(emb)) in one launch
        t = self._timestep_tensor(timestep, B)
        emb = o.timestep_sinusoid(t, self.cfg.block_out_channels[0])
        emb = o.small_linear(emb, P["t1w"], P["t1b"], silu_out=True)
        emb = o.small_linear(emb, P["t2w"], P["t2b"])
        st.temb = o.small_linear(emb, P["temb_w"], P["temb_b"], silu_in=True)

        x = o.conv_in(sample.float().contiguous(), P["cin_w"], P["cin_b"])
        skips = [x]
        for blk in self.down_blocks:
            x, s, h, w = blk.run(st, x, h, w)
            skips += s
        x = self.mid_block.run(st, x, h, w)
        for blk in self.up_blocks:
            x, h, w = blk.run(st, x, skips, h, w)
        a = o.group_norm(x, None, nimg, h * w, self.conv_norm_out.num_groups, P["gn_g"], P["gn_b"],
                         self.conv_norm_out.eps, silu=True)
        out = o.conv_out(a, B, F, h, w, P["cout_w"], P["cout_b"])
        if sample.dtype != torch.float32:
            out = out.to(sample.dtype)
        return UNetMotionOutput(sample=out)

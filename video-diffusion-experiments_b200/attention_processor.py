"""Attention-processor surface (diffusers: ``processor(attn, hidden_states, encoder_hidden_states=None,
attention_mask=None, temb=None, **kw) -> Tensor``; north star: "keep the ... attention-processor ... surface").

``B200AttnProcessor`` lets any module with ``to_q / to_k / to_v / to_out[0] / heads`` (our ``Attention`` or a stock
diffusers ``Attention``) run its projections and softmax(QK^T)V on the b200ad kernels for a stand-alone
``[Bt, L, C]`` call.  Inside ``UNetMotionModel.forward`` the fused plan bypasses processors (the same kernels are
called with fused epilogues and without the layout round trip).
"""
from __future__ import annotations

import torch

from .ops import get_ops


def _head_pad(w: torch.Tensor, heads: int, dpad: int) -> torch.Tensor:
    n, k = w.shape
    d = n // heads
    if d == dpad:
        return w
    out = w.new_zeros(heads, dpad, k)
    out[:, :d] = w.view(heads, d, k)
    return out.view(heads * dpad, k)


class B200AttnProcessor:
    def __init__(self, mode: str = "bf16"):
        self.mode = mode

    @torch.no_grad()
    def __call__(self, attn, hidden_states, encoder_hidden_states=None, attention_mask=None, temb=None, **kw):
        if attention_mask is not None:
            raise NotImplementedError("B200AttnProcessor: attention masks are not on the AnimateDiff path")
        if not hidden_states.is_cuda:
            raise RuntimeError("B200AttnProcessor needs CUDA tensors (no CPU fallback exists)")
        o = get_ops(getattr(attn, "_b200_mode", self.mode))
        x = hidden_states
        ctx = x if encoder_hidden_states is None else encoder_hidden_states
        bt, lq, c = x.shape
        lk = ctx.shape[1]
        heads = attn.heads
        d = attn.to_q.weight.shape[0] // heads
        dpad = d if o.mode == "fp32" else next(p for p in (64, 128, 192) if d <= p)
        dev = lambda w: w.detach().to("cuda", o.op_dtype).contiguous()
        xr = o.cast(x.reshape(bt * lq, c).float().contiguous(), 1, 1, bt * lq, c)
        cr = xr if encoder_hidden_states is None else o.cast(
            ctx.reshape(-1, ctx.shape[-1]).float().contiguous(), 1, 1, ctx.shape[0] * lk, ctx.shape[-1])
        _, q = o.gemm(xr, dev(_head_pad(attn.to_q.weight.detach(), heads, dpad)), want_op=True)
        _, k = o.gemm(cr, dev(_head_pad(attn.to_k.weight.detach(), heads, dpad)), want_op=True)
        _, v = o.gemm(cr, dev(_head_pad(attn.to_v.weight.detach(), heads, dpad)), want_op=True)
        kv_div = bt // ctx.shape[0]
        att = o.attention(q, k, v, bt, heads, lq, lk, d, dpad, kv_div)
        lin = attn.to_out[0]
        bias = None if lin.bias is None else lin.bias.detach().to("cuda", torch.float32).contiguous()
        out, _ = o.gemm(att, dev(lin.weight), bias=bias, want_f32=True)
        return out.view(bt, lq, -1).to(hidden_states.dtype)

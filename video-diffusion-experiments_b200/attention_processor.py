"""Attention-processor surface (diffusers: ``processor(attn, hidden_states, encoder_hidden_states=None,
attention_mask=None, temb=None, **kw) -> Tensor``; north star: "keep the ... attention-processor ... surface").

``B200AttnProcessor`` is this repo's AttnProcessor2_0: it runs ``attn.to_q / to_k / to_v`` (called as modules, so
forward hooks on them fire -- /root/reference/utils/forward_tracer.py:177-194), softmax(Q K^T / sqrt(d)) V on the
b200ad attention kernels, then ``attn.to_out[0]`` / ``attn.to_out[1]``.  It serves the stand-alone ``[Bt, L, C]`` call
of a product ``Attention`` (module mode) and, given a stock diffusers ``Attention`` whose projections are plain
``torch.nn.Linear``, runs those projections on b200ad_gemm directly from the module's weights.  Inside the fused plan
(``UNetMotionModel.forward`` without hooks) processors are bypassed: the same kernels run with fused epilogues.
"""
from __future__ import annotations

import torch

from .leaf import Linear as _B200Linear
from .ops import get_ops


def _proj(o, lin, x):
    """lin(x) for [Bt, L, K]: through the module when it is a b200ad leaf (hooks fire), else on b200ad_gemm from the
    module's weights (a stock torch.nn.Linear would run cuBLAS)."""
    if isinstance(lin, _B200Linear):
        return lin(x)
    bt, l, k = x.shape
    w = lin.weight.detach().to("cuda", o.op_dtype).contiguous()
    b = None if lin.bias is None else lin.bias.detach().to("cuda", torch.float32).contiguous()
    y, _ = o.gemm(o.cast(x.reshape(bt * l, k).float().contiguous(), 1, 1, bt * l, k), w, bias=b, want_f32=True)
    return y.view(bt, l, -1).to(x.dtype)


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
        bt, lq, _ = x.shape
        lk = ctx.shape[1]
        heads = attn.heads
        q, k, v = _proj(o, attn.to_q, x), _proj(o, attn.to_k, ctx), _proj(o, attn.to_v, ctx)
        c = q.shape[-1]
        d = c // heads
        if encoder_hidden_states is None and lq <= 32 and d % 8 == 0:
            # short sequences, many of them (the motion module's [B*h*w, F, C]): the frame-axis kernel, one warp per
            # (sequence, head); rows are (sequence, position)-ordered, i.e. its (b, f, hw) layout with hw = 1
            qkv = torch.cat([q, k, v], -1).reshape(bt * lq, 3 * c).float().contiguous()
            att = o.temporal_attention(o.cast(qkv, 1, 1, bt * lq, 3 * c), bt, lq, 1, c, heads)
        else:
            dpad = d if o.mode == "fp32" else next(p for p in (64, 128, 192) if d <= p)

            def heads_padded(t, rows):
                buf = torch.zeros(rows, heads, dpad, device=t.device, dtype=torch.float32)
                buf[..., :d] = t.reshape(rows, heads, d).float()
                return o.cast(buf.view(rows, heads * dpad), 1, 1, rows, heads * dpad)

            kv_div = bt // ctx.shape[0]
            att = o.attention(heads_padded(q, bt * lq), heads_padded(k, ctx.shape[0] * lk),
                              heads_padded(v, ctx.shape[0] * lk), bt, heads, lq, lk, d, dpad, kv_div)
        out = att.view(bt, lq, c).to(x.dtype)
        out = _proj(o, attn.to_out[0], out)
        return attn.to_out[1](out) if len(attn.to_out) > 1 else out


AttnProcessor2_0 = B200AttnProcessor

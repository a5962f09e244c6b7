"""AutoencoderKL *decoder* of SD1.5 on the b200ad sm_100a kernels -- row f1 of SURVEY.md section 8:
``decode_latents`` + ``postprocess_video``, the step right after the denoise loop whose frames the reference consumes
at /root/reference/experiments/05_grid_search_ablation.py:169,178-179.

Module tree and state-dict keys follow diffusers (``post_quant_conv``, ``decoder.conv_in``, ``decoder.mid_block.
{resnets.N, attentions.0.{group_norm,to_q,to_k,to_v,to_out.0}}``, ``decoder.up_blocks.i.{resnets.j,upsamplers.0.conv}``,
``decoder.conv_norm_out``, ``decoder.conv_out``), the leaves are the b200ad leaf modules (leaf.py), and ``decode``
runs a fused plan on channels-last row matrices like the UNet:

  latents/0.18215 -> post_quant_conv (1x1) -> conv_in : ONE direct kernel (b200ad_conv_in_affine; the 1x1 conv and the
      scaling are folded into the 3x3 weights, its bias rides on a ones channel so the zero padding stays exact)
  ResnetBlock2D : GroupNorm+SiLU (b200ad_group_norm) -> tcgen05 implicit-GEMM conv3x3 (+bias) -> GroupNorm+SiLU ->
      conv3x3 (+bias, +fp32 residual / 1x1 shortcut GEMM) -- the UNet's kernels at C = 512/256/128, up to 512^2 pixels
  mid attention (1 head, d = 512, L = h*w): GroupNorm -> Q, K GEMMs -> per image: S = Q K^T (tcgen05 GEMM, fp32 out) ->
      b200ad_softmax_rows -> O = P V (GEMM against V^T, itself produced by a GEMM with swapped operands; V's bias is
      folded into the out-projection bias since softmax rows sum to 1) -> out-projection GEMM (+bias, +residual)
  Upsample2D : nearest x2 written once in the operand type (b200ad_cast mode 2) -> conv3x3
  conv_norm_out+SiLU -> conv_out (128 -> 3, warp-MMA kernel writing [1,3,F,H,W]) -> b200ad_video_to_u8 (x/2+.5, clamp,
      *255, round) for the uint8 frames.

No CPU fallback, no cuDNN/cuBLAS: every computation is a libb200ad kernel.
"""
from __future__ import annotations

import os
from typing import Optional, Sequence

import torch
import torch.nn as nn

from .leaf import Conv2d, GroupNorm, Linear, SiLU
from .ops import Ops, _empty, get_ops
from .unet import _Packable, _dev_f, _dev_w, _mode_of, _pack_conv3x3, _pack_upconv3x3


class _VaeResnet(_Packable):
    """diffusers ResnetBlock2D with temb_channels=None (GroupNorm eps 1e-6)."""

    def __init__(self, cin: int, cout: int, groups: int = 32):
        super().__init__()
        self.in_channels, self.out_channels = cin, cout
        self.norm1 = GroupNorm(groups, cin, eps=1e-6)
        self.conv1 = Conv2d(cin, cout, 3, padding=1)
        self.norm2 = GroupNorm(groups, cout, eps=1e-6)
        self.dropout = nn.Dropout(0.0)
        self.conv2 = Conv2d(cout, cout, 3, padding=1)
        self.nonlinearity = SiLU()
        self.conv_shortcut = Conv2d(cin, cout, 1) if cin != cout else None

    def _pack(self, o: Ops):
        p = {"g1": _dev_f(self.norm1.weight), "b1": _dev_f(self.norm1.bias),
             "g2": _dev_f(self.norm2.weight), "b2": _dev_f(self.norm2.bias),
             "w1": _pack_conv3x3(o, self.conv1.weight), "cb1": _dev_f(self.conv1.bias),
             "w2": _pack_conv3x3(o, self.conv2.weight), "cb2": _dev_f(self.conv2.bias)}
        if self.conv_shortcut is not None:
            w = self.conv_shortcut.weight
            p["wsc"], p["bsc"] = _dev_w(o, w.reshape(w.shape[0], w.shape[1])), _dev_f(self.conv_shortcut.bias)
            if o.mode == "bf16" and self.in_channels % 64 == 0 and os.environ.get("B200AD_CONV_SC", "1") != "0":
                p["w2sc"] = torch.cat([p["w2"], p["wsc"]], 1).contiguous()      # see unet.ResnetBlock2D._pack
                p["b2sc"] = p["cb2"] + p["bsc"]
        return p

    def run(self, o: Ops, x, n: int, h: int, w: int):
        """x: fp32 rows [n*h*w, cin] -> fp32 rows [n*h*w, cout]."""
        p = self.packed(o)
        G = self.norm1.num_groups
        need_raw = self.conv_shortcut is not None
        a = o.group_norm(x, None, n, h * w, G, p["g1"], p["b1"], self.norm1.eps, silu=True, want_raw=need_raw)
        raw = None
        if need_raw:
            a, raw = a
        h32, _ = o.conv3x3(a, p["w1"], n, h, w, bias=p["cb1"], stats=True)
        del a
        a2 = o.group_norm(h32, None, n, h * w, G, p["g2"], p["b2"], self.norm2.eps, silu=True)
        del h32
        if need_raw and "w2sc" in p:
            return o.conv3x3_sc(a2, p["w2sc"], raw, n, h, w, bias=p["b2sc"], stats=True)
        res = o.gemm(raw, p["wsc"], bias=p["bsc"], want_f32=True)[0] if need_raw else x
        out, _ = o.conv3x3(a2, p["w2"], n, h, w, bias=p["cb2"], residual=res, stats=True)
        return out

    def forward(self, input_tensor, temb=None, *a, **kw):
        """Module mode: every sub-module called (hooks fire), NCHW."""
        from . import leaf as _leaf
        h = self.conv1(self.nonlinearity(self.norm1(input_tensor)))
        h = self.conv2(self.dropout(self.nonlinearity(self.norm2(h))))
        x = input_tensor if self.conv_shortcut is None else self.conv_shortcut(input_tensor)
        return _leaf.add(get_ops(_mode_of(self)), x, h)


class _VaeAttention(_Packable):
    """diffusers Attention(heads=1, dim_head=C, bias=True, norm_num_groups=32, residual_connection=True)."""

    def __init__(self, c: int, groups: int = 32):
        super().__init__()
        self.heads = 1
        self.group_norm = GroupNorm(groups, c, eps=1e-6)
        self.to_q, self.to_k, self.to_v = Linear(c, c), Linear(c, c), Linear(c, c)
        self.to_out = nn.ModuleList([Linear(c, c), nn.Dropout(0.0)])

    def _pack(self, o: Ops):
        wo, bo = self.to_out[0].weight.detach().float(), self.to_out[0].bias.detach().float()
        bv = self.to_v.bias.detach().float()
        return {"gn_g": _dev_f(self.group_norm.weight), "gn_b": _dev_f(self.group_norm.bias),
                "wq": _dev_w(o, self.to_q.weight), "bq": _dev_f(self.to_q.bias),
                # softmax is invariant to a per-row constant: q.(k + bk) = q.k + q.bk, so K's bias drops out exactly
                "wk": _dev_w(o, self.to_k.weight),
                "wv": _dev_w(o, self.to_v.weight),
                "wo": _dev_w(o, wo),
                # softmax rows sum to 1: P (V + 1 bv^T) Wo^T + bo = (P V) Wo^T + (Wo bv + bo)
                "bo": _dev_f(bo + wo @ bv)}

    def run(self, o: Ops, x, n: int, hw: int):
        """x: fp32 rows [n*hw, C], updated in place (x += attention(GN(x)))."""
        p = self.packed(o)
        C = x.shape[1]
        a = o.group_norm(x, None, n, hw, self.group_norm.num_groups, p["gn_g"], p["gn_b"], self.group_norm.eps,
                         silu=False)
        _, q = o.gemm(a, p["wq"], bias=p["bq"], want_op=True)                        # [n*hw, C]
        _, k = o.gemm(a, p["wk"], want_op=True)
        att = _empty((n * hw, C), x.device, o.op_dtype)
        s32 = _empty((hw, hw), x.device, torch.float32)
        prob = _empty((hw, hw), x.device, o.op_dtype)
        vt = _empty((C, hw), x.device, o.op_dtype)
        scale = float(C) ** -0.5
        for i in range(n):
            rows = slice(i * hw, (i + 1) * hw)
            o.gemm(p["wv"], a[rows], out_op=vt)                                       # V^T = Wv A^T   [C, hw]
            o.gemm(q[rows], k[rows], out_f32=s32)                                     # S = Q K^T      [hw, hw]
            o.softmax_rows(s32, scale, out=prob)
            o.gemm(prob, vt, out_op=att[rows])                                        # O = P V        [hw, C]
        o.gemm(att, p["wo"], bias=p["bo"], residual=x, out_f32=x, stats_hw=hw)        # x feeds the next resnet's GroupNorm
        return x

    def forward(self, hidden_states, *a, **kw):
        """Module mode: NCHW in/out through to_q / to_k / to_v / to_out (hooks fire); softmax(QK^T/sqrt(C))V by GEMMs."""
        from . import leaf as _leaf
        o = get_ops(_mode_of(self))
        n, c, h, w = hidden_states.shape
        t = self.group_norm(hidden_states).reshape(n, c, h * w).transpose(1, 2)       # [n, hw, c]
        q, k, v = self.to_q(t), self.to_k(t), self.to_v(t)
        outs = []
        for i in range(n):
            qi = o.cast(q[i].float().contiguous(), 1, 1, h * w, c)
            ki = o.cast(k[i].float().contiguous(), 1, 1, h * w, c)
            vti = o.cast(v[i].float().t().contiguous(), 1, 1, c, h * w)
            s32, _ = o.gemm(qi, ki, want_f32=True)
            outs.append(o.gemm(o.softmax_rows(s32, float(c) ** -0.5), vti, want_f32=True)[0])
        att = torch.stack(outs).to(hidden_states.dtype)
        out = self.to_out[1](self.to_out[0](att)).transpose(1, 2).reshape(n, c, h, w).contiguous()
        return _leaf.add(o, out, hidden_states)


class _VaeMid(nn.Module):
    def __init__(self, c: int, groups: int = 32):
        super().__init__()
        self.attentions = nn.ModuleList([_VaeAttention(c, groups)])
        self.resnets = nn.ModuleList([_VaeResnet(c, c, groups), _VaeResnet(c, c, groups)])

    def forward(self, x, *a, **kw):
        return self.resnets[1](self.attentions[0](self.resnets[0](x)))


class _VaeUpsample(_Packable):
    def __init__(self, c: int):
        super().__init__()
        self.conv = Conv2d(c, c, 3, padding=1)

    def _pack(self, o: Ops):
        sub = o.mode == "bf16" and os.environ.get("B200AD_UPCONV", "1") != "0"     # see unet.Upsample2D._pack
        return {"w": None if sub else _pack_conv3x3(o, self.conv.weight), "b": _dev_f(self.conv.bias),
                "w4": _pack_upconv3x3(o, self.conv.weight) if sub else None}

    def run(self, o: Ops, x, n: int, h: int, w: int):
        p = self.packed(o)
        if p["w4"] is not None:
            return o.upconv3x3(o.cast(x, n, h, w, x.shape[1]), p["w4"], n, h, w, bias=p["b"], stats=True)
        up = o.cast(x, n, h, w, x.shape[1], mode=2)
        return o.conv3x3(up, p["w"], n, 2 * h, 2 * w, bias=p["b"], stats=True)[0]

    def forward(self, x, *a, **kw):
        from . import leaf as _leaf
        n, c, h, w = x.shape
        up = get_ops("fp32").cast(_leaf.nchw_to_rows(x), n, h, w, c, mode=2)
        return self.conv(_leaf.rows_to_nchw(up, n, 2 * h, 2 * w, x.dtype))


class _VaeUpBlock(nn.Module):
    def __init__(self, cin: int, cout: int, up: bool, groups: int = 32):
        super().__init__()
        self.resnets = nn.ModuleList([_VaeResnet(cin if j == 0 else cout, cout, groups) for j in range(3)])
        self.upsamplers = nn.ModuleList([_VaeUpsample(cout)]) if up else None

    def forward(self, x, *a, **kw):
        for r in self.resnets:
            x = r(x)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class _Decoder(nn.Module):
    def __init__(self, chans: Sequence[int] = (512, 512, 256, 128), groups: int = 32):
        super().__init__()
        self.conv_in = Conv2d(4, chans[0], 3, padding=1)
        self.mid_block = _VaeMid(chans[0], groups)
        self.up_blocks = nn.ModuleList()
        prev = chans[0]
        for i, c in enumerate(chans):
            self.up_blocks.append(_VaeUpBlock(prev, c, up=i < len(chans) - 1, groups=groups))
            prev = c
        self.conv_norm_out = GroupNorm(groups, prev, eps=1e-6)
        self.conv_act = SiLU()
        self.conv_out = Conv2d(prev, 3, 3, padding=1)

    def forward(self, z, *a, **kw):
        """Module mode (hooks fire on every sub-module)."""
        x = self.mid_block(self.conv_in(z))
        for b in self.up_blocks:
            x = b(x)
        return self.conv_out(self.conv_act(self.conv_norm_out(x)))


class AutoencoderKLDecoder(_Packable):
    """``decode_video(latents)``: [1,4,F,h,w] latents (as the denoise loop leaves them) -> [1,3,F,8h,8w] fp32 video;
    ``decode_video_u8``: the same through postprocess_video -> uint8 frames [F,8h,8w,3];
    ``decode(z)``: diffusers' AutoencoderKL.decode on [N,4,h,w] (already divided by the scaling factor)."""
    scaling_factor = 0.18215

    def __init__(self, chans: Sequence[int] = (512, 512, 256, 128), groups: int = 32, mode: str = "bf16"):
        super().__init__()
        self.post_quant_conv = Conv2d(4, 4, 1)
        self.decoder = _Decoder(chans, groups)
        self.use_slicing = False
        self.frames_per_pass = 16
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
    def mode(self) -> str:
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
        d = self.decoder
        win = d.conv_in.weight.detach().float().cpu()                           # [C0, 4, 3, 3]
        wpq = self.post_quant_conv.weight.detach().float().cpu()[:, :, 0, 0]     # [4(j), 4(c)]
        bpq = self.post_quant_conv.bias.detach().float().cpu()
        p = {"cin_b": _dev_f(d.conv_in.bias), "gn_g": _dev_f(d.conv_norm_out.weight),
             "gn_b": _dev_f(d.conv_norm_out.bias), "cout_w": _dev_f(d.conv_out.weight),
             "cout_b": _dev_f(d.conv_out.bias)}
        for name, s in (("cin_w_lat", self.scaling_factor), ("cin_w_z", 1.0)):
            w = torch.einsum("ojyx,jc->ocyx", win.double(), wpq.double() / s)
            ones = torch.einsum("ojyx,j->oyx", win.double(), bpq.double())[:, None]
            p[name] = _dev_f(torch.cat([w, ones], 1).float())                    # [C0, 5, 3, 3]
        return p

    # ---- fused plan --------------------------------------------------------------------------------
    @torch.no_grad()
    def _decode_rows(self, x5, key: str):
        """x5: fp32 [1, 4, n, h, w] -> fp32 [1, 3, n, 8h, 8w]."""
        if not x5.is_cuda:
            raise RuntimeError("AutoencoderKLDecoder: CUDA tensors required (the b200ad path has no CPU fallback)")
        o = get_ops(self.mode)
        o.colstats.clear()
        P = self.packed(o)
        d = self.decoder
        _, _, n, h, w = x5.shape
        x = o.conv_in_affine(x5.float().contiguous(), P[key], P["cin_b"])
        x = d.mid_block.resnets[0].run(o, x, n, h, w)
        x = d.mid_block.attentions[0].run(o, x, n, h * w)
        x = d.mid_block.resnets[1].run(o, x, n, h, w)
        for blk in d.up_blocks:
            for r in blk.resnets:
                x = r.run(o, x, n, h, w)
            if blk.upsamplers is not None:
                x = blk.upsamplers[0].run(o, x, n, h, w)
                h, w = 2 * h, 2 * w
        a = o.group_norm(x, None, n, h * w, d.conv_norm_out.num_groups, P["gn_g"], P["gn_b"], d.conv_norm_out.eps,
                         silu=True)
        del x
        return o.conv_out(a, 1, n, h, w, P["cout_w"], P["cout_b"])

    def _hooks_registered(self) -> bool:
        return any(m is not self and (m._forward_hooks or m._forward_pre_hooks) for m in self.modules())

    @torch.no_grad()
    def decode_video(self, latents: torch.Tensor) -> torch.Tensor:
        """[1,4,F,h,w] latents -> [1,3,F,8h,8w] fp32 (decode_latents of AnimateDiffPipeline: /scaling_factor inside)."""
        if latents.dim() != 5 or latents.shape[0] != 1:
            raise ValueError(f"decode_video expects [1,4,F,h,w], got {tuple(latents.shape)}")
        F = latents.shape[2]
        step = 1 if self.use_slicing else self.frames_per_pass
        if F <= step:
            return self._decode_rows(latents, "cin_w_lat")
        return torch.cat([self._decode_rows(latents[:, :, i:i + step], "cin_w_lat") for i in range(0, F, step)], 2)

    @torch.no_grad()
    def decode_video_u8(self, latents: torch.Tensor) -> torch.Tensor:
        """Latents -> uint8 frames [F, 8h, 8w, 3] (decode + postprocess_video on the device)."""
        v = self.decode_video(latents)
        return get_ops(self.mode).video_to_u8(v[0].contiguous())

    @torch.no_grad()
    def decode(self, z: torch.Tensor, return_dict: bool = False):
        """diffusers AutoencoderKL.decode: z [N,4,h,w] (already / scaling_factor) -> [N,3,8h,8w]."""
        if self._hooks_registered():
            img = self.decoder(self.post_quant_conv(z))
        else:
            x5 = z.permute(1, 0, 2, 3)[None]                                     # [1, 4, N, h, w]
            N = z.shape[0]
            step = 1 if self.use_slicing else self.frames_per_pass
            out = torch.cat([self._decode_rows(x5[:, :, i:i + step], "cin_w_z") for i in range(0, N, step)], 2)
            img = out[0].permute(1, 0, 2, 3).to(z.dtype)
        return DecoderOutput(sample=img) if return_dict else img

    def forward(self, z, *a, **kw):
        return self.decode(z)


class DecoderOutput:
    def __init__(self, sample):
        self.sample = sample

"""Leaf modules of the diffusers module tree on the b200ad kernels -- the hook-visible "module mode".

The fused denoise step (``UNetMotionModel.forward`` -> ``.run()`` methods) never calls a sub-module: it launches
fused kernels on channels-last row matrices.  The reference, however, also *inspects* the model:
/root/reference/utils/forward_tracer.py:177-194 registers a forward hook on every named module and
/root/reference/experiments/03_trace_forward_pass.py:105-113,182-205 expects them to fire with diffusers' tensor
shapes (``[B*H*W, F, C]`` at the temporal attention, 03:160-169).  For that the tree is built from these leaves:
``Linear`` / ``Conv2d`` / ``GroupNorm`` / ``LayerNorm`` / ``SiLU`` are ``torch.nn`` subclasses with the same names,
parameters and state-dict keys whose ``forward`` runs the corresponding libb200ad kernel *unfused* (one module, one
call), in NCHW / ``[B, L, C]`` tensors like diffusers.  Same kernels, same operand rounding points as the fused plan
(operands are rounded to bf16 where the fused plan rounds them); only the fusion is given up.

No CPU fallback: every forward needs CUDA tensors.  Layout changes (permute / reshape copies) are torch views and
copies; all arithmetic is a libb200ad kernel.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .ops import Ops, get_ops


def mode_of(m: nn.Module) -> str:
    return getattr(m, "_b200_mode", "bf16")


def _need_cuda(x: torch.Tensor):
    if not x.is_cuda:
        raise RuntimeError("b200ad modules need CUDA tensors (no CPU fallback exists)")


def rows_f32(x: torch.Tensor) -> torch.Tensor:
    """[..., C] -> contiguous fp32 [rows, C]."""
    return x.reshape(-1, x.shape[-1]).float().contiguous()


def nchw_to_rows(x: torch.Tensor) -> torch.Tensor:
    n, c, h, w = x.shape
    return x.float().permute(0, 2, 3, 1).reshape(n * h * w, c).contiguous()


def rows_to_nchw(r: torch.Tensor, n: int, h: int, w: int, dtype) -> torch.Tensor:
    return r.view(n, h, w, -1).permute(0, 3, 1, 2).to(dtype)


class _LeafCache:
    """Operand-typed copies of a leaf's parameters, rebuilt when the parameter storage or the mode changes."""

    def _packed(self, o: Ops, build):
        key = (o.mode,) + tuple((p.data_ptr(), p._version) for p in self.parameters(recurse=False))
        c = self.__dict__.get("_b200_leaf")
        if c is None or c[0] != key:
            with torch.no_grad():
                c = (key, build())
            self.__dict__["_b200_leaf"] = c
        return c[1]


class Linear(nn.Linear, _LeafCache):
    """``y = x W^T + b`` on b200ad_gemm (tcgen05, bf16 operands; fp32 SIMT in the fp32 kernel mode).  Up to 64 rows
    (the time-embedding MLP and ``time_emb_proj``) run on the fp32 small-M kernel, as in the fused plan."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _need_cuda(x)
        o = get_ops(mode_of(self))
        r = rows_f32(x)
        M, K = r.shape
        if M <= 64 and K % 4 == 0:
            p = self._packed(get_ops("fp32"), lambda: {
                "w": self.weight.detach().float().contiguous(),
                "b": None if self.bias is None else self.bias.detach().float().contiguous()})
            y = torch.cat([o.small_linear(r[i:i + 8].contiguous(), p["w"], p["b"]) for i in range(0, M, 8)], 0)
        else:
            p = self._packed(o, lambda: {
                "w": self.weight.detach().to(o.op_dtype).contiguous(),
                "b": None if self.bias is None else self.bias.detach().float().contiguous()})
            y, _ = o.gemm(o.cast(r, 1, 1, M, K), p["w"], bias=p["b"], want_f32=True)
        return y.view(*x.shape[:-1], self.out_features).to(x.dtype)


class Conv2d(nn.Conv2d, _LeafCache):
    """1x1 (GEMM over pixel rows), 3x3 pad 1 stride 1 / stride 2 (tcgen05 implicit GEMM), and the UNet's 4-channel
    ends: cin <= 8 runs conv_in's direct kernel, cout <= 4 runs conv_out's."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _need_cuda(x)
        o = get_ops(mode_of(self))
        n, cin, h, w = x.shape
        cout = self.out_channels
        k, s, pad = self.kernel_size[0], self.stride[0], self.padding[0]
        bias = lambda: None if self.bias is None else self.bias.detach().float().contiguous()
        if k == 1 and s == 1 and pad == 0:
            # the tensor-core GEMM wants N % 32 == 0 and K % 8 == 0: tiny 1x1 convs (the VAE's post_quant_conv, 4 -> 4)
            # are zero-padded to that shape
            kp, np_ = (cin + 7) // 8 * 8, (cout + 31) // 32 * 32

            def build():
                wp = torch.zeros(np_, kp, device=self.weight.device, dtype=torch.float32)
                wp[:cout, :cin] = self.weight.detach().reshape(cout, cin).float()
                bp = None
                if self.bias is not None:
                    bp = torch.zeros(np_, device=self.weight.device, dtype=torch.float32)
                    bp[:cout] = self.bias.detach().float()
                return {"w": wp.to(o.op_dtype).contiguous(), "b": bp}
            p = self._packed(o, build)
            r = nchw_to_rows(x)
            if kp != cin:
                r = torch.nn.functional.pad(r, (0, kp - cin))
            y, _ = o.gemm(o.cast(r, 1, 1, r.shape[0], kp), p["w"], bias=p["b"], want_f32=True)
            if np_ != cout:
                y = y[:, :cout].contiguous()
            return rows_to_nchw(y, n, h, w, x.dtype)
        if not (k == 3 and pad == 1 and s in (1, 2)):
            raise NotImplementedError(f"b200ad Conv2d: kernel {k} stride {s} padding {pad} is not on the AnimateDiff path")
        if cin <= 8 and s == 1:                 # conv_in: reads [B, cin, F, h, w] with F = 1
            p = self._packed(get_ops("fp32"), lambda: {"w": self.weight.detach().float().contiguous(), "b": bias()})
            y = o.conv_in(x.float().contiguous().view(n, cin, 1, h, w), p["w"], p["b"])
            return rows_to_nchw(y, n, h, w, x.dtype)
        if cout <= 4 and s == 1:                # conv_out: writes [B, cout, F, h, w] with F = 1
            p = self._packed(get_ops("fp32"), lambda: {"w": self.weight.detach().float().contiguous(), "b": bias()})
            r = nchw_to_rows(x)
            y = o.conv_out(o.cast(r, 1, 1, r.shape[0], cin), n, 1, h, w, p["w"], p["b"])
            return y.view(n, cout, h, w).to(x.dtype)
        p = self._packed(o, lambda: {
            "w": self.weight.detach().permute(0, 2, 3, 1).reshape(cout, -1).to(o.op_dtype).contiguous(), "b": bias()})
        r = nchw_to_rows(x)
        if s == 1:
            y, _ = o.conv3x3(o.cast(r, n, h, w, cin), p["w"], n, h, w, bias=p["b"])
            return rows_to_nchw(y, n, h, w, x.dtype)
        y, _ = o.conv3x3(o.cast(r, n, h, w, cin, mode=1), p["w"], n, h // 2, w // 2, stride2=True, bias=p["b"])
        return rows_to_nchw(y, n, h // 2, w // 2, x.dtype)


class GroupNorm(nn.GroupNorm, _LeafCache):
    """GroupNorm on b200ad_group_norm.  4-D ``[N, C, h, w]``: per-image statistics; 5-D ``[B, C, F, h, w]`` (the
    motion module): statistics over all F frames of a batch item (frames_per_stat = F)."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _need_cuda(x)
        o = get_ops("fp32")        # fp32 output: the consumer (Conv2d / Linear leaf) rounds once to its operand type
        p = self._packed(o, lambda: {"g": self.weight.detach().float().contiguous(),
                                     "b": self.bias.detach().float().contiguous()})
        if x.dim() == 4:
            n, c, h, w = x.shape
            y = o.group_norm(nchw_to_rows(x), None, n, h * w, self.num_groups, p["g"], p["b"], self.eps, silu=False)
            return rows_to_nchw(y, n, h, w, x.dtype)
        if x.dim() == 5:
            b, c, f, h, w = x.shape
            r = x.float().permute(0, 2, 3, 4, 1).reshape(b * f * h * w, c).contiguous()
            y = o.group_norm(r, None, b * f, h * w, self.num_groups, p["g"], p["b"], self.eps, silu=False,
                             frames_per_stat=f)
            return y.view(b, f, h, w, c).permute(0, 4, 1, 2, 3).to(x.dtype)
        raise NotImplementedError(f"b200ad GroupNorm: {x.dim()}-D input")


class LayerNorm(nn.LayerNorm, _LeafCache):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _need_cuda(x)
        o = get_ops("fp32")        # fp32 output (see GroupNorm)
        p = self._packed(o, lambda: {"g": self.weight.detach().float().contiguous(),
                                     "b": self.bias.detach().float().contiguous()})
        y = o.layer_norm(rows_f32(x), p["g"], p["b"], self.eps)
        return y.view(x.shape).to(x.dtype)


class SiLU(nn.SiLU):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _need_cuda(x)
        return get_ops(mode_of(self)).silu(x.float().contiguous()).view(x.shape).to(x.dtype)


def add(o: Ops, a: torch.Tensor, b: torch.Tensor, div: int = 1, mod: Optional[int] = None) -> torch.Tensor:
    """a + b (same shape), or a + b broadcast as b.flat[(i // div) % mod] (fp32 kernel; result in a's dtype)."""
    af, bf_ = a.float().contiguous(), b.float().contiguous()
    return o.add(af, bf_, div=div, mod=mod).view(a.shape).to(a.dtype)

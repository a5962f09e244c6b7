"""Functional wrappers over the C ABI: torch tensors in, torch tensors out.

PyTorch is plumbing here (device memory, the current stream); every computation is a kernel in libb200ad.so.
All functions require CUDA tensors and raise otherwise -- there is no CPU fallback.
"""
from __future__ import annotations

import os
import weakref
from typing import Optional

import torch

from . import lib as _lib

BF16, FP32 = 0, 1

# Diagnostic: POISON = True fills every buffer this module allocates with NaN (0xFF for uint8) instead of leaving it
# uninitialised, so a kernel that reads bytes no kernel has written yet turns the result into NaN instead of silently
# picking up whatever the caching allocator last kept there (tests/test_unet_gpu.py::test_no_read_before_write).
POISON = os.environ.get("B200AD_POISON") == "1"


def _empty(shape, device, dtype):
    if POISON:
        if dtype in (torch.uint8, torch.int32, torch.int64):
            return torch.full(tuple(shape), 255 if dtype == torch.uint8 else -1, device=device, dtype=dtype)
        return torch.full(tuple(shape), float("nan"), device=device, dtype=dtype)
    return torch.empty(tuple(shape), device=device, dtype=dtype)


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.B200ADError("b200ad kernels need CUDA tensors (no CPU fallback exists)")


class KernelProfiler:
    """Per-kernel-family timing with CUDA events on the launching stream (eager launches only; bench.py uses it
    for the roofline numbers).  Each record: family, algorithmic FLOPs, algorithmic bytes, (start, end) events."""

    def __init__(self):
        self.records = []

    def summary(self):
        torch.cuda.synchronize()
        fam = {}
        for name, flops, nbytes, e0, e1, launches in self.records:
            d = fam.setdefault(name, {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
            d["launches"] += launches
            d["ms"] += e0.elapsed_time(e1)
            d["flops"] += flops
            d["bytes"] += nbytes
        for d in fam.values():
            sec = max(d["ms"], 1e-9) * 1e-3
            d["tflops"] = d["flops"] / sec / 1e12
            d["gbs"] = d["bytes"] / sec / 1e9
        return fam


class _Rec:
    __slots__ = ("prof", "name", "flops", "nbytes", "e0", "launches")

    def __init__(self, prof, name, flops, nbytes, launches=1):
        self.prof, self.name, self.flops, self.nbytes, self.launches = prof, name, flops, nbytes, launches

    def __enter__(self):
        if self.prof is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *exc):
        if self.prof is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            self.prof.records.append((self.name, self.flops, self.nbytes, self.e0, e1, self.launches))
        return False


def _nb(*ts):
    return float(sum(t.numel() * t.element_size() for t in ts if t is not None))


class Ops:
    """Kernel front-end bound to one operand mode: 'bf16' (tcgen05 tensor cores) or 'fp32' (SIMT)."""
    profiler: Optional[KernelProfiler] = None       # class-level: set by bench.py around an eager pass

    def _rec(self, name, flops=0.0, nbytes=0.0, launches=1):
        return _Rec(Ops.profiler, name, flops, nbytes, launches)

    # GroupNorm column statistics emitted by GEMM / conv epilogues: data_ptr of the fp32 output -> (stats, spi, nimg,
    # hw, C, out tensor).  A later writer of the same buffer without statistics drops the entry.
    use_epilogue_stats = os.environ.get("B200AD_EPILOGUE_STATS", "1") != "0"      # A/B switch for measurements

    def __init__(self, mode: str = "bf16"):
        if mode not in ("bf16", "fp32"):
            raise ValueError(f"mode must be 'bf16' or 'fp32', got {mode!r}")
        self.colstats = {}
        self.mode = mode
        self.op = BF16 if mode == "bf16" else FP32
        self.op_dtype = torch.bfloat16 if mode == "bf16" else torch.float32
        self.L = _lib.load()

    # ---------------------------------------------------------------- contractions
    def _stats_ok(self, conv: bool, h: int, w: int) -> int:
        """32-row statistic slots per image the epilogue can emit for this geometry, or 0 (fp32 mode / unsupported)."""
        if self.op != BF16 or not Ops.use_epilogue_stats:
            return 0
        spi = self.L.b200ad_stats_slots(1 if conv else 0, h, w)
        return spi if spi > 0 else 0

    def _forget_stats(self, out_f32):
        if out_f32 is not None and self.colstats:
            self.colstats.pop(out_f32.data_ptr(), None)

    def gemm(self, a, w, *, bias=None, row_bias=None, rb_div=1, residual=None, out_f32=None, out_op=None,
             want_f32=False, want_op=False, geglu=False, stats_hw=None):
        """C = a[M,K] @ w[N,K]^T with fused epilogue (b200ad_gemm).  Returns (out_f32, out_op).
        stats_hw (rows per image): the fp32 output feeds a GroupNorm -- its per-column sums are emitted by the epilogue
        (b200ad_gemm_stats) and remembered for group_norm()."""
        _need_cuda(a, w)
        assert a.dim() == 2 and w.dim() == 2 and a.dtype == self.op_dtype and w.dtype == self.op_dtype
        assert a.stride(1) == 1 and w.is_contiguous()
        M, K = a.shape
        N = w.shape[0]
        assert w.shape[1] == K
        n_out = N // 2 if geglu else N
        if want_f32 and out_f32 is None:
            out_f32 = _empty((M, n_out), a.device, torch.float32)
        if want_op and out_op is None:
            out_op = _empty((M, n_out), a.device, self.op_dtype)
        spi = self._stats_ok(False, stats_hw, 1) if (stats_hw and out_f32 is not None and out_op is None and not geglu
                                                     and M % stats_hw == 0) else 0
        if spi:
            nimg = M // stats_hw
            st = _empty((nimg * spi, N, 2), a.device, torch.float32)
            with self._rec("gemm", 2.0 * M * N * K, _nb(w, residual, out_f32, st) + M * K * a.element_size()):
                rc = self.L.b200ad_gemm_stats(
                    self.op, a.data_ptr(), a.stride(0), w.data_ptr(), M, N, K, _ptr(bias), _ptr(row_bias), rb_div,
                    0 if row_bias is None else row_bias.stride(0), _ptr(residual),
                    0 if residual is None else residual.stride(0), out_f32.data_ptr(), out_f32.stride(0),
                    st.data_ptr(), stats_hw, _stream())
            _lib.check(rc, "b200ad_gemm_stats")
            self.colstats[out_f32.data_ptr()] = (st, spi, nimg, stats_hw, N, weakref.ref(out_f32))
            return out_f32, None
        self._forget_stats(out_f32)
        with self._rec("gemm", 2.0 * M * N * K, _nb(w, residual, out_f32, out_op) + M * K * a.element_size()):
            rc = self.L.b200ad_gemm(
                self.op, a.data_ptr(), a.stride(0), w.data_ptr(), M, N, K, _ptr(bias), _ptr(row_bias), rb_div,
                0 if row_bias is None else row_bias.stride(0), _ptr(residual),
                0 if residual is None else residual.stride(0), _ptr(out_f32),
                0 if out_f32 is None else out_f32.stride(0), _ptr(out_op),
                0 if out_op is None else out_op.stride(0), 1 if geglu else 0, _stream())
        _lib.check(rc, "b200ad_gemm")
        return out_f32, out_op

    def conv3x3(self, src, w, nimg, h, wd, *, stride2=False, bias=None, row_bias=None, rb_div=1, residual=None,
                want_f32=True, want_op=False, stats=False):
        """3x3 conv over NHWC `src` (operand-typed); w [cout, 9*cin].  Returns (out_f32 [nimg*h*wd, cout], out_op).
        stats: the fp32 output feeds a GroupNorm (see gemm)."""
        _need_cuda(src, w)
        assert src.dtype == self.op_dtype and w.dtype == self.op_dtype and src.is_contiguous() and w.is_contiguous()
        cout, k9 = w.shape
        cin = k9 // 9
        M = nimg * h * wd
        out_f32 = _empty((M, cout), src.device, torch.float32) if want_f32 else None
        out_op = _empty((M, cout), src.device, self.op_dtype) if want_op else None
        spi = self._stats_ok(True, h, wd) if (stats and want_f32 and not want_op) else 0
        if spi:
            st = _empty((nimg * spi, cout, 2), src.device, torch.float32)
            with self._rec("conv3x3", 2.0 * M * cout * k9, _nb(src, w, residual, out_f32, st)):
                rc = self.L.b200ad_conv3x3_stats(
                    self.op, src.data_ptr(), w.data_ptr(), nimg, h, wd, cin, cout, 1 if stride2 else 0, _ptr(bias),
                    _ptr(row_bias), rb_div, 0 if row_bias is None else row_bias.stride(0), _ptr(residual),
                    0 if residual is None else residual.stride(0), out_f32.data_ptr(), cout, st.data_ptr(), _stream())
            _lib.check(rc, "b200ad_conv3x3_stats")
            self.colstats[out_f32.data_ptr()] = (st, spi, nimg, h * wd, cout, weakref.ref(out_f32))
            return out_f32, None
        with self._rec("conv3x3", 2.0 * M * cout * k9, _nb(src, w, residual, out_f32, out_op)):
            rc = self.L.b200ad_conv3x3(
                self.op, src.data_ptr(), w.data_ptr(), nimg, h, wd, cin, cout, 1 if stride2 else 0, _ptr(bias),
                _ptr(row_bias), rb_div, 0 if row_bias is None else row_bias.stride(0), _ptr(residual),
                0 if residual is None else residual.stride(0), _ptr(out_f32), cout, _ptr(out_op), cout, _stream())
        _lib.check(rc, "b200ad_conv3x3")
        return out_f32, out_op

    def conv3x3_sc(self, src, w, sc_src, nimg, h, wd, *, bias=None, stats=False):
        """conv3x3(src) + 1x1 conv of sc_src accumulated in the same tiles (b200ad_conv3x3_sc): w [cout, 9*cin + sc_cin],
        src [nimg*h*wd, cin], sc_src [nimg*h*wd, sc_cin], both operand-typed.  Returns the fp32 rows.  bf16 mode only."""
        _need_cuda(src, w, sc_src)
        assert self.op == BF16, "conv3x3_sc exists on the tcgen05 path only"
        assert src.dtype == self.op_dtype and w.dtype == self.op_dtype and sc_src.dtype == self.op_dtype
        assert src.is_contiguous() and w.is_contiguous() and sc_src.is_contiguous()
        cout, ktot = w.shape
        cin, sc_cin = src.shape[1], sc_src.shape[1]
        assert ktot == 9 * cin + sc_cin
        M = nimg * h * wd
        out = _empty((M, cout), src.device, torch.float32)
        spi = self._stats_ok(True, h, wd) if stats else 0
        st = _empty((nimg * spi, cout, 2), src.device, torch.float32) if spi else None
        with self._rec("conv3x3", 2.0 * M * cout * ktot, _nb(src, sc_src, w, out, st)):
            rc = self.L.b200ad_conv3x3_sc(self.op, src.data_ptr(), w.data_ptr(), nimg, h, wd, cin, cout,
                                          sc_src.data_ptr(), sc_cin, _ptr(bias), out.data_ptr(), cout, None, 0,
                                          _ptr(st), _stream())
        _lib.check(rc, "b200ad_conv3x3_sc")
        if spi:
            self.colstats[out.data_ptr()] = (st, spi, nimg, h * wd, cout, weakref.ref(out))
        return out

    def upconv3x3(self, src, w4, nimg, h, wd, *, bias=None, stats=False):
        """nearest-x2 upsample + 3x3 conv of the low-resolution NHWC `src` [nimg*h*wd, cin] (operand-typed) as four
        2x2 sub-pixel convolutions (b200ad_upconv3x3); w4 [4, cout, 4*cin] from pack_upconv_weights().  Returns the
        fp32 rows of the [nimg, 2h, 2wd] image.  bf16 mode only.  FLOPs booked are the executed ones (4/9 of the
        direct form's)."""
        _need_cuda(src, w4)
        assert self.op == BF16, "upconv3x3 exists on the tcgen05 path only"
        assert src.dtype == self.op_dtype and w4.dtype == self.op_dtype and src.is_contiguous() and w4.is_contiguous()
        _, cout, k4 = w4.shape
        cin = k4 // 4
        M = nimg * 4 * h * wd
        out = _empty((M, cout), src.device, torch.float32)
        spi = self._stats_ok(True, h, wd) if stats else 0
        st = _empty((nimg * 4 * spi, cout, 2), src.device, torch.float32) if spi else None
        with self._rec("conv3x3", 2.0 * M * cout * k4, _nb(src, w4, out, st)):
            rc = self.L.b200ad_upconv3x3(self.op, src.data_ptr(), w4.data_ptr(), nimg, h, wd, cin, cout, _ptr(bias),
                                         out.data_ptr(), cout, None, 0, _ptr(st), _stream())
        _lib.check(rc, "b200ad_upconv3x3")
        if spi:
            self.colstats[out.data_ptr()] = (st, 4 * spi, nimg, 4 * h * wd, cout, weakref.ref(out))
        return out

    def attention(self, q, k, v, nbatch, heads, Lq, Lk, d, dpad, kv_div=1, v_ones=False):
        """q/k/v: 2-D operand-typed views (row stride = leading dim).  Returns [nbatch*Lq, heads*d].
        v_ones: pad column d of every head of v holds 1.0 (B200AD_ATTN_V_ONES)."""
        _need_cuda(q, k, v)
        out = _empty((nbatch * Lq, heads * d), q.device, self.op_dtype)
        es = out.element_size()
        nbytes = es * heads * d * (2.0 * nbatch * Lq + 2.0 * (nbatch // kv_div) * Lk)
        with self._rec("attention_L%d_d%d" % (Lk, d), 4.0 * nbatch * heads * Lq * Lk * d, nbytes):
            rc = self.L.b200ad_attention2(self.op, q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(),
                                          v.stride(0), nbatch, heads, Lq, Lk, d, dpad, kv_div, 1 if v_ones else 0,
                                          out.data_ptr(), out.stride(0), _stream())
        _lib.check(rc, "b200ad_attention2")
        return out

    def temporal_attention(self, qkv, B, F, hw, C, heads):
        _need_cuda(qkv)
        assert qkv.is_contiguous() and qkv.shape == (B * F * hw, 3 * C)
        out = _empty((B * F * hw, C), qkv.device, self.op_dtype)
        with self._rec("temporal_attention", 4.0 * B * hw * F * F * C, _nb(qkv, out)):
            rc = self.L.b200ad_temporal_attention(self.op, qkv.data_ptr(), B, F, hw, C, heads, out.data_ptr(),
                                                  _stream())
        _lib.check(rc, "b200ad_temporal_attention")
        return out

    # ---------------------------------------------------------------- norms / re-layout
    def group_norm(self, x0, x1, nimg, hw, groups, gamma, beta, eps, *, silu, frames_per_stat=1, want_raw=False):
        """GroupNorm(+SiLU) of cat([x0, x1], channel) for fp32 [nimg*hw, c] inputs -> operand-typed output."""
        _need_cuda(x0, x1)
        assert x0.dtype == torch.float32 and x0.is_contiguous()
        c0 = x0.shape[-1]
        c1 = 0 if x1 is None else x1.shape[-1]
        if x1 is not None:
            assert x1.dtype == torch.float32 and x1.is_contiguous()
        out = _empty((nimg * hw, c0 + c1), x0.device, self.op_dtype)
        raw = _empty(out.shape, out.device, out.dtype) if want_raw else None
        # statistics from the producers' epilogues (every source must have them, for this geometry)
        s0 = self._lookup_stats(x0, nimg, hw, c0)
        s1 = self._lookup_stats(x1, nimg, hw, c1) if x1 is not None else None
        if s0 is not None and (x1 is None or s1 is not None) and (c0 + c1) // groups % 2 == 0 and c0 % 2 == 0:
            ws = _empty((int(self.L.b200ad_gn_stats_ws_bytes(nimg, groups, frames_per_stat)) // 8,), x0.device, torch.float64)
            # algorithmic bytes: the fp32 input ONCE + the operand-typed output(s)
            with self._rec("group_norm", 0.0, _nb(x0, x1) + _nb(out, raw)):
                rc = self.L.b200ad_gn_apply_stats(
                    self.op, x0.data_ptr(), c0, s0[0].data_ptr(), s0[1], _ptr(x1), c1,
                    None if s1 is None else s1[0].data_ptr(), 0 if s1 is None else s1[1], nimg, hw, groups,
                    frames_per_stat, float(eps), gamma.data_ptr(), beta.data_ptr(), 1 if silu else 0, ws.data_ptr(),
                    out.data_ptr(), _ptr(raw), _stream())
            _lib.check(rc, "b200ad_gn_apply_stats")
            return (out, raw) if want_raw else out
        nblk = self.L.b200ad_gn_nblk(hw)
        partial = _empty((nimg * nblk * groups + nimg * groups, 2), x0.device, torch.float32)
        # algorithmic bytes (1 read + 1 write is the floor): the fp32 input once + the operand-typed output(s); this
        # path READS the input twice (statistics pass, apply pass)
        with self._rec("group_norm", 0.0, _nb(x0, x1) + _nb(out, raw)):
            rc = self.L.b200ad_group_norm(self.op, x0.data_ptr(), c0, _ptr(x1), c1, nimg, hw, groups,
                                          partial.data_ptr(), frames_per_stat, float(eps), gamma.data_ptr(),
                                          beta.data_ptr(), 1 if silu else 0, out.data_ptr(), _ptr(raw), _stream())
        _lib.check(rc, "b200ad_group_norm")
        return (out, raw) if want_raw else out

    def _lookup_stats(self, x, nimg, hw, c):
        e = self.colstats.get(x.data_ptr()) if self.colstats else None
        if e is None or e[5]() is not x or e[2:5] != (nimg, hw, c):
            return None
        return e

    def layer_norm(self, x, gamma, beta, eps=1e-5, pe=None, hw=1, frames=1):
        _need_cuda(x)
        assert x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 2
        rows, C = x.shape
        out = _empty((rows, C), x.device, self.op_dtype)
        with self._rec("layernorm", 0.0, _nb(x, out)):
            rc = self.L.b200ad_layernorm(self.op, x.data_ptr(), rows, C, gamma.data_ptr(), beta.data_ptr(),
                                         float(eps), _ptr(pe), hw, frames, out.data_ptr(), _stream())
        _lib.check(rc, "b200ad_layernorm")
        return out

    def cast(self, x, nimg, h, w, C, mode=0):
        """fp32 NHWC -> operand type; mode 0 plain, 1 space-to-depth phase split, 2 nearest x2 upsample."""
        _need_cuda(x)
        assert x.dtype == torch.float32 and x.is_contiguous()
        if mode == 0:
            shape = (nimg * h * w, C)
        elif mode == 1:
            shape = (4 * nimg * (h // 2) * (w // 2), C)
        else:
            shape = (nimg * 4 * h * w, C)
        out = _empty(shape, x.device, self.op_dtype)
        with self._rec("cast", 0.0, _nb(out) * (1.0 + 4.0 / out.element_size() / (4 if mode == 2 else 1))):
            rc = self.L.b200ad_cast(self.op, x.data_ptr(), nimg, h, w, C, mode, out.data_ptr(), _stream())
        _lib.check(rc, "b200ad_cast")
        return out

    # ---------------------------------------------------------------- small ops
    def timestep_sinusoid(self, t, dim):
        _need_cuda(t)
        assert t.dtype == torch.float32 and t.dim() == 1
        out = _empty((t.shape[0], dim), t.device, torch.float32)
        _lib.check(self.L.b200ad_timestep_sinusoid(t.data_ptr(), t.shape[0], dim, out.data_ptr(), _stream()),
                   "b200ad_timestep_sinusoid")
        return out

    def small_linear(self, x, w, b, silu_in=False, silu_out=False):
        _need_cuda(x, w)
        assert x.dtype == torch.float32 and w.dtype == torch.float32 and x.is_contiguous() and w.is_contiguous()
        M, K = x.shape
        N = w.shape[0]
        y = _empty((M, N), x.device, torch.float32)
        with self._rec("small_linear", 2.0 * M * N * K, _nb(w, x, y)):
            _lib.check(self.L.b200ad_small_linear(x.data_ptr(), M, K, w.data_ptr(), _ptr(b), N, int(silu_in),
                                                  int(silu_out), y.data_ptr(), _stream()), "b200ad_small_linear")
        return y

    def conv_in(self, sample, w, b):
        _need_cuda(sample, w)
        assert sample.dtype == torch.float32 and sample.is_contiguous() and sample.dim() == 5
        B, cin, F, h, wd = sample.shape
        cout = w.shape[0]
        out = _empty((B * F * h * wd, cout), sample.device, torch.float32)
        with self._rec("conv_in", 2.0 * B * F * h * wd * cout * cin * 9, _nb(sample, out)):
            _lib.check(self.L.b200ad_conv_in(sample.data_ptr(), B, cin, F, h, wd, w.data_ptr(), _ptr(b), cout,
                                             out.data_ptr(), _stream()), "b200ad_conv_in")
        return out

    def conv_out(self, act, B, F, h, wd, w, b):
        _need_cuda(act, w)
        cout, cin = w.shape[0], w.shape[1]
        assert act.dtype == self.op_dtype and act.is_contiguous()
        out = _empty((B, cout, F, h, wd), act.device, torch.float32)
        with self._rec("conv_out", 2.0 * B * F * h * wd * cout * cin * 9, _nb(act, out)):
            _lib.check(self.L.b200ad_conv_out(self.op, act.data_ptr(), B, F, h, wd, cin, w.data_ptr(), _ptr(b), cout,
                                              out.data_ptr(), _stream()), "b200ad_conv_out")
        return out

    # ---------------------------------------------------------------- module-mode glue / VAE pieces
    def _eltwise(self, mode, a, b, n, div, mod, out):
        _lib.check(self.L.b200ad_eltwise(mode, a.data_ptr(), _ptr(b), n, div, mod, out.data_ptr(), _stream()),
                   "b200ad_eltwise")
        return out

    def silu(self, x):
        _need_cuda(x)
        assert x.dtype == torch.float32 and x.is_contiguous()
        with self._rec("eltwise", 0.0, 2 * _nb(x)):
            return self._eltwise(0, x, None, x.numel(), 1, 1, _empty(x.shape, x.device, x.dtype))

    def add(self, a, b, div: int = 1, mod=None):
        """a + b.flat[(i // div) % mod] (mod defaults to b.numel(): plain add when b has a's size)."""
        _need_cuda(a, b)
        assert a.dtype == torch.float32 and b.dtype == torch.float32 and a.is_contiguous() and b.is_contiguous()
        mod = b.numel() if mod is None else mod
        assert mod <= b.numel() and (div != 1 or mod != b.numel() or b.numel() <= a.numel())
        with self._rec("eltwise", 0.0, 2 * _nb(a) + _nb(b)):
            return self._eltwise(1, a, b, a.numel(), div, mod, _empty(a.shape, a.device, a.dtype))

    def quick_gelu(self, x):
        _need_cuda(x)
        assert x.dtype == torch.float32 and x.is_contiguous()
        with self._rec("eltwise", 0.0, 2 * _nb(x)):
            return self._eltwise(3, x, None, x.numel(), 1, 1, _empty(x.shape, x.device, x.dtype))

    def embed_tokens(self, ids, tok_table, pos_table):
        """ids int64 [B, seq] -> fp32 [B*seq, dim] = token_table[ids] + position_table[:seq]."""
        _need_cuda(ids, tok_table, pos_table)
        assert ids.dtype == torch.int64 and ids.is_contiguous() and tok_table.is_contiguous() and pos_table.is_contiguous()
        B, seq = ids.shape
        dim = tok_table.shape[1]
        out = _empty((B * seq, dim), ids.device, torch.float32)
        _lib.check(self.L.b200ad_embed_tokens(ids.data_ptr(), B * seq, seq, tok_table.shape[0], tok_table.data_ptr(),
                                              pos_table.data_ptr(), dim, out.data_ptr(), _stream()), "b200ad_embed_tokens")
        return out

    def attention_small(self, q, k, v, nbatch, heads, L, d, causal=False):
        """Short-sequence attention (optionally causal) on 2-D operand-typed views -> [nbatch*L, heads*d]."""
        _need_cuda(q, k, v)
        out = _empty((nbatch * L, heads * d), q.device, self.op_dtype)
        with self._rec("attention_small", 4.0 * nbatch * heads * L * L * d, _nb(out) * 4):
            _lib.check(self.L.b200ad_attention_small(self.op, q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0),
                                                     v.data_ptr(), v.stride(0), nbatch, heads, L, d, 1 if causal else 0,
                                                     out.data_ptr(), out.stride(0), _stream()), "b200ad_attention_small")
        return out

    def geglu_gate(self, hg):
        """[rows, 2C] = (h | g) -> [rows, C] = h * gelu_erf(g)."""
        _need_cuda(hg)
        assert hg.dtype == torch.float32 and hg.is_contiguous() and hg.dim() == 2 and hg.shape[1] % 2 == 0
        rows, c2 = hg.shape
        out = _empty((rows, c2 // 2), hg.device, torch.float32)
        with self._rec("eltwise", 0.0, _nb(hg, out)):
            return self._eltwise(2, hg, None, out.numel(), 1, c2 // 2, out)

    def softmax_rows(self, s, scale: float, out=None):
        """softmax(scale * s) along the last axis of fp32 [rows, cols] (row stride = s.stride(0)) -> operand type."""
        _need_cuda(s)
        assert s.dtype == torch.float32 and s.dim() == 2 and s.stride(1) == 1
        rows, cols = s.shape
        if out is None:
            out = _empty((rows, cols), s.device, self.op_dtype)
        with self._rec("softmax_rows", 0.0, rows * cols * (4.0 + out.element_size())):
            _lib.check(self.L.b200ad_softmax_rows(self.op, s.data_ptr(), rows, cols, s.stride(0), float(scale),
                                                  out.data_ptr(), out.stride(0), _stream()), "b200ad_softmax_rows")
        return out

    def video_to_u8(self, video):
        """[C, F, h, w] fp32 in [-1, 1] -> [F, h, w, C] uint8 (diffusers postprocess_video)."""
        _need_cuda(video)
        assert video.dtype == torch.float32 and video.is_contiguous() and video.dim() == 4
        C, F, h, w = video.shape
        out = _empty((F, h, w, C), video.device, torch.uint8)
        with self._rec("video_to_u8", 0.0, _nb(video, out)):
            _lib.check(self.L.b200ad_video_to_u8(video.data_ptr(), C, F, h, w, out.data_ptr(), _stream()),
                       "b200ad_video_to_u8")
        return out

    def conv_in_affine(self, sample, w_aug, b):
        """conv_in over [sample | ones] (b200ad_conv_in_affine): sample [B, cin, F, h, w], w_aug [cout, cin+1, 3, 3]."""
        _need_cuda(sample, w_aug)
        assert sample.dtype == torch.float32 and sample.is_contiguous() and sample.dim() == 5
        assert w_aug.dtype == torch.float32 and w_aug.is_contiguous()
        B, cin, F, h, wd = sample.shape
        cout = w_aug.shape[0]
        assert w_aug.shape[1] == cin + 1
        out = _empty((B * F * h * wd, cout), sample.device, torch.float32)
        with self._rec("conv_in", 2.0 * B * F * h * wd * cout * (cin + 1) * 9, _nb(sample, out)):
            _lib.check(self.L.b200ad_conv_in_affine(sample.data_ptr(), B, cin, F, h, wd, w_aug.data_ptr(), _ptr(b), cout,
                                                    out.data_ptr(), _stream()), "b200ad_conv_in_affine")
        return out

    def cfg_step(self, eps_u, eps_c, latents, g, cx, ce, in_scale=1.0, latents_out=None, model_in=None):
        """latents' = cx*latents + ce*(eps_u + g*(eps_c-eps_u)); model_in[k] = latents'*in_scale."""
        _need_cuda(eps_u, eps_c, latents)
        for t in (eps_u, eps_c, latents):
            assert t.dtype == torch.float32 and t.is_contiguous()
        n = latents.numel()
        if latents_out is None:
            latents_out = _empty(latents.shape, latents.device, latents.dtype)
        copies = 0 if model_in is None else model_in.numel() // n
        _lib.check(self.L.b200ad_cfg_step(eps_u.data_ptr(), eps_c.data_ptr(), latents.data_ptr(), float(g),
                                          float(cx), float(ce), float(in_scale), n, latents_out.data_ptr(),
                                          _ptr(model_in), copies, _stream()), "b200ad_cfg_step")
        return latents_out


    def cfg_step_dev(self, eps_u, eps_c, latents, coef4, latents_out, model_in=None):
        """cfg_step with (g, cx, ce, in_scale) in device memory `coef4` (graph-replayable)."""
        _need_cuda(eps_u, eps_c, latents, coef4)
        n = latents.numel()
        copies = 0 if model_in is None else model_in.numel() // n
        with self._rec("cfg_step", 0.0, 4.0 * n * (4 + copies)):
            _lib.check(self.L.b200ad_cfg_step_dev(eps_u.data_ptr(), eps_c.data_ptr(), latents.data_ptr(),
                                                  coef4.data_ptr(), n, latents_out.data_ptr(), _ptr(model_in),
                                                  copies, _stream()), "b200ad_cfg_step_dev")
        return latents_out


_OPS = {}


def get_ops(mode: str = "bf16") -> Ops:
    if mode not in _OPS:
        _OPS[mode] = Ops(mode)
    return _OPS[mode]

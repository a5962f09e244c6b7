#!/usr/bin/env python
"""Per-kernel microbenchmarks at the bench workload's real shapes (CUDA events, L2 flushed between runs):
    python tools/kbench.py temporal|attention|gemm|conv|norm|all
Prints one line per shape: avg us, achieved TFLOP/s or GB/s and the fraction of the measured peak."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import b200ad  # noqa: E402,F401
from b200ad.ops import get_ops  # noqa: E402

try:
    PK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    PK = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
o = get_ops("bf16")
FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    tot = 0.0
    for _ in range(iters):
        FLUSH.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3   # us


def report(name, us, flops=0.0, nbytes=0.0):
    tf, gb = flops / us / 1e6, nbytes / us / 1e3
    print(f"{name:58s} {us:9.1f} us  {tf:8.1f} TF/s ({tf / PK['bf16_tflops']:.2f})  {gb:8.1f} GB/s ({gb / PK['hbm_gbs']:.2f})",
          flush=True)


def bf(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(torch.bfloat16)


def temporal():
    for hw, C in [(4096, 320), (1024, 640), (256, 1280), (64, 1280)]:
        qkv = bf(2 * 16 * hw, 3 * C)
        us = timeit(lambda: o.temporal_attention(qkv, 2, 16, hw, C, 8))
        report(f"temporal_attention B=2 F=16 hw={hw} C={C}", us, 4.0 * 2 * hw * 16 * 16 * C, qkv.numel() * 2 * 4 / 3)


def attention():
    for L, C, dp in [(4096, 320, 64), (1024, 640, 128), (256, 1280, 192)]:
        d = C // 8
        qkv = bf(32 * L, 3 * 8 * dp)
        hd = 8 * dp
        us = timeit(lambda: o.attention(qkv[:, :hd], qkv[:, hd:2 * hd], qkv[:, 2 * hd:], 32, 8, L, L, d, dp, 1), 5)
        report(f"self-attention n=32 heads=8 L={L} d={d} (dpad {dp})", us, 4.0 * 32 * 8 * L * L * d, 32 * L * C * 2 * 4)
        if dp > d and L >= 512:
            qkv.view(32 * L, 3, 8, dp)[:, 2, :, d] = 1.0            # V pad column d = 1 (B200AD_ATTN_V_ONES)
            us = timeit(lambda: o.attention(qkv[:, :hd], qkv[:, hd:2 * hd], qkv[:, 2 * hd:], 32, 8, L, L, d, dp, 1,
                                            v_ones=True), 5)
            report(f"   same, denominator from the PV product (V ones column)", us, 4.0 * 32 * 8 * L * L * d,
                   32 * L * C * 2 * 4)
        kv = bf(2 * 77, 2 * hd)
        q = bf(32 * L, hd)
        us = timeit(lambda: o.attention(q, kv[:, :hd], kv[:, hd:], 32, 8, L, 77, d, dp, 16), 5)
        report(f"cross-attention n=32 heads=8 Lq={L} Lk=77 d={d}", us, 4.0 * 32 * 8 * L * 77 * d, 32 * L * C * 2 * 2)


def gemm():
    shapes = []
    for hw, C in [(4096, 320), (1024, 640), (256, 1280), (64, 1280)]:
        M = 32 * hw
        shapes += [(M, C, C, "res32"), (M, 3 * C, C, "op"), (M, 8 * C, C, "geglu"), (M, C, 4 * C, "res32"),
                   (M, C, 4 * C, "resop"), (M, C, 5 * C, "res32")]     # K = 5C: the folded FF2 + proj_out GEMM
    for M, N, K, kind in shapes:
        a, w = bf(M, K), bf(N, K, scale=K ** -0.5)
        bias = torch.randn(N, device="cuda")
        res = torch.randn(M, N, device="cuda") if kind.startswith("res") else None
        if kind == "res32":
            fn = lambda: o.gemm(a, w, bias=bias, residual=res, out_f32=res)
            nb = M * K * 2 + N * K * 2 + 8.0 * M * N
        elif kind == "resop":
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            fn = lambda: o.gemm(a, w, bias=bias, residual=res, out_op=out)
            nb = M * K * 2 + N * K * 2 + 6.0 * M * N
        elif kind == "op":
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            fn = lambda: o.gemm(a, w, out_op=out)
            nb = M * K * 2 + N * K * 2 + 2.0 * M * N
        else:
            out = torch.empty(M, N // 2, device="cuda", dtype=torch.bfloat16)
            fn = lambda: o.gemm(a, w, bias=bias, out_op=out, geglu=True)
            nb = M * K * 2 + N * K * 2 + 1.0 * M * N
        us = timeit(fn, 5)
        report(f"gemm M={M} N={N} K={K} {kind}", us, 2.0 * M * N * K, nb)


def conv():
    for n, cin, cout, r in [(32, 320, 320, 64), (32, 640, 320, 64), (32, 640, 640, 32), (32, 1280, 1280, 16),
                            (32, 1280, 1280, 8), (32, 2560, 1280, 16), (32, 2560, 1280, 8), (32, 1920, 640, 32)]:
        src = bf(n * r * r, cin)
        w = bf(cout, 9 * cin, scale=(9 * cin) ** -0.5)
        bias = torch.randn(cout, device="cuda")
        us = timeit(lambda: o.conv3x3(src, w, n, r, r, bias=bias), 5)
        report(f"conv3x3 n={n} {cin}->{cout} @{r}", us, 2.0 * n * r * r * cout * 9 * cin,
               src.numel() * 2 + w.numel() * 2 + 4.0 * n * r * r * cout)


def norm():
    for hw, C in [(4096, 320), (1024, 640), (256, 1280)]:
        x = torch.randn(32 * hw, C, device="cuda")
        g, b = torch.randn(C, device="cuda"), torch.randn(C, device="cuda")
        us = timeit(lambda: o.layer_norm(x, g, b))
        report(f"layernorm rows={32 * hw} C={C}", us, 0, x.numel() * 6.0)
        o.colstats.clear()
        us = timeit(lambda: o.group_norm(x, None, 32, hw, 32, g, b, 1e-5, silu=True))
        report(f"groupnorm(stats+apply) n=32 hw={hw} C={C}", us, 0, x.numel() * 6.0)
        # statistics from the producer's epilogue: a GEMM writes x (+ column sums), then finalize + apply only
        a, w = bf(32 * hw, 64), bf(C, 64, scale=0.125)
        o.gemm(a, w, residual=x, out_f32=x, stats_hw=hw)
        us = timeit(lambda: o.group_norm(x, None, 32, hw, 32, g, b, 1e-5, silu=True))
        report(f"groupnorm(epilogue stats: finalize+apply) n=32 hw={hw} C={C}", us, 0, x.numel() * 6.0)
        us0 = timeit(lambda: o.gemm(a, w, residual=x, out_f32=x))
        us1 = timeit(lambda: o.gemm(a, w, residual=x, out_f32=x, stats_hw=hw))
        print(f"   producer GEMM (K=64) without / with column statistics: {us0:.1f} / {us1:.1f} us", flush=True)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    for name in ("temporal", "attention", "gemm", "conv", "norm"):
        if which in (name, "all"):
            globals()[name]()

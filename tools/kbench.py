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


def restructured():
    """The restructured convolutions of DESIGN 3.11 against the module-order forms they replace."""
    from b200ad.unet import _pack_upconv3x3
    # Upsample2D of the UNet's up blocks 0 / 1 / 2 (source resolution 8 / 16 / 32 at 16f 512^2)
    for n, c, r in [(32, 1280, 8), (32, 1280, 16), (32, 640, 32)]:
        x = torch.randn(n * r * r, c, device="cuda")
        wt = torch.randn(c, c, 3, 3) * (9 * c) ** -0.5
        w9 = wt.permute(0, 2, 3, 1).reshape(c, -1).to("cuda", torch.bfloat16).contiguous()
        w4 = _pack_upconv3x3(o, wt)
        bias = torch.randn(c, device="cuda")
        us0 = timeit(lambda: o.conv3x3(o.cast(x, n, r, r, c, mode=2), w9, n, 2 * r, 2 * r, bias=bias, stats=True), 5)
        us1 = timeit(lambda: o.upconv3x3(o.cast(x, n, r, r, c), w4, n, r, r, bias=bias, stats=True), 5)
        fl = 2.0 * n * 4 * r * r * c * 9 * c
        report(f"upsample x2 + conv3x3 n={n} C={c} @{r}->{2 * r}: cast(mode 2) + conv", us0, fl, 0)
        report(f"    sub-pixel: cast + 4 x (2x2 conv) [4/9 of the MACs]", us1, fl * 4 / 9, 0)
    # ResnetBlock2D.conv2 + conv_shortcut of the up blocks (cin = skip concat)
    for n, cin, cout, r in [(32, 960, 320, 64), (32, 640, 320, 64), (32, 1920, 640, 32), (32, 2560, 1280, 16),
                            (32, 2560, 1280, 8)]:
        a2, raw = bf(n * r * r, cout), bf(n * r * r, cin)
        w2, wsc = bf(cout, 9 * cout, scale=(9 * cout) ** -0.5), bf(cout, cin, scale=cin ** -0.5)
        wcat = torch.cat([w2, wsc], 1).contiguous()
        bias = torch.randn(cout, device="cuda")

        def two():
            res, _ = o.gemm(raw, wsc, bias=bias, want_f32=True)
            o.conv3x3(a2, w2, n, r, r, bias=bias, residual=res, stats=True)
        us0 = timeit(two, 5)
        us1 = timeit(lambda: o.conv3x3_sc(a2, wcat, raw, n, r, r, bias=bias, stats=True), 5)
        fl = 2.0 * n * r * r * cout * (9 * cout + cin)
        report(f"conv2 + shortcut n={n} {cout}(+{cin} sc) @{r}: gemm + conv", us0, fl, 0)
        report(f"    one launch (second K segment)", us1, fl, 0)
    # last block's FF2 + proj_out: two GEMMs against the folded K = 5C GEMM (+ the operand copy in attn2's out-projection)
    for hw, C in [(4096, 320), (1024, 640), (256, 1280), (64, 1280)]:
        M = 32 * hw
        g, h, x = bf(M, 4 * C), torch.randn(M, C, device="cuda"), torch.randn(M, C, device="cuda")
        w2, wo = bf(C, 4 * C, scale=(4 * C) ** -0.5), bf(C, C, scale=C ** -0.5)
        cat, wcat = bf(M, 5 * C), bf(C, 5 * C, scale=(5 * C) ** -0.5)
        bias = torch.randn(C, device="cuda")

        def two():
            hb = o.gemm(g, w2, bias=bias, residual=h, want_op=True)[1]
            o.gemm(hb, wo, bias=bias, residual=x, out_f32=x, stats_hw=hw)
        us0 = timeit(two, 5)
        us1 = timeit(lambda: o.gemm(cat, wcat, bias=bias, residual=x, out_f32=x, stats_hw=hw), 5)
        fl = 2.0 * M * C * 5 * C
        report(f"FF2 + proj_out M={M} C={C}: two GEMMs", us0, fl, 0)
        report(f"    folded K = 5C GEMM", us1, fl, 0)


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
    for name in ("temporal", "attention", "gemm", "conv", "norm", "restructured"):
        if which in (name, "all"):
            globals()[name]()

#!/usr/bin/env python
"""Where the tcgen05 GEMM's producer and UMMA-issuer threads wait (cycle counters written by the kernel itself):
    python tools/gemm_waits.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import b200ad  # noqa
from b200ad import lib
from b200ad.ops import get_ops
o = get_ops("bf16")
L = lib.load()
dbg = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")

def run(name, fn):
    for _ in range(2): fn()
    L.b200ad_debug_buffer(dbg.data_ptr()); dbg.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    L.b200ad_debug_buffer(None)
    d = dbg.view(148, 16).double().cpu()
    lead, peer = d[0::2], d[1::2]
    print(f"{name}: {e0.elapsed_time(e1)*1e3:.1f} us | producer(leader) total {lead[:,0].mean():.0f} wait_empty {lead[:,1].mean():.0f} | "
          f"producer(peer) total {peer[:,0].mean():.0f} wait_empty {peer[:,1].mean():.0f} | mma total {lead[:,2].mean():.0f} "
          f"wait_full {lead[:,3].mean():.0f} wait_acc {lead[:,4].mean():.0f} | W-producer total {lead[:,5].mean():.0f} wait_empty {lead[:,6].mean():.0f}\n      epilogue(g0 lead warp): total {lead[:,8].mean():.0f} wait_tfull {lead[:,9].mean():.0f} wait_buf {lead[:,10].mean():.0f} | io warp total {lead[:,7].mean():.0f} wait_chunk {lead[:,11].mean():.0f} issue {lead[:,15].mean():.0f} wait_read {lead[:,12].mean():.0f} | bias_stage {lead[:,13].mean():.0f} chunks {lead[:,14].mean():.0f}", flush=True)

def bf(*s, scale=1.0): return (torch.randn(*s, device="cuda") * scale).to(torch.bfloat16)
which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "big"):
    for (M, N, K) in [(8192, 8192, 8192), (131072, 960, 320), (8192, 10240, 1280)]:
        a, w = bf(M, K), bf(N, K, scale=K ** -0.5)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        run(f"gemm {M}x{N}x{K} op", lambda: o.gemm(a, w, out_op=out))
        print(f"   -> {2.0*M*N*K/1e12:.2f} TFLOP")
    for n, cin, cout, r in [(32, 320, 320, 64), (32, 1280, 1280, 16)]:
        src, w = bf(n * r * r, cin), bf(cout, 9 * cin, scale=(9 * cin) ** -0.5)
        run(f"conv {cin}->{cout}@{r}", lambda: o.conv3x3(src, w, n, r, r))
if which in ("small",):
    # the deepest level (M = 2048 rows, one tile per cluster) and level 2 (M = 8192)
    for (M, N, K, kind) in [(2048, 1280, 1280, "res32"), (2048, 1280, 5120, "res32"), (2048, 3840, 1280, "op"),
                            (2048, 10240, 1280, "geglu"), (8192, 1280, 1280, "res32"), (8192, 1280, 5120, "res32")]:
        a, w = bf(M, K), bf(N, K, scale=K ** -0.5)
        bias = torch.randn(N, device="cuda")
        if kind == "op":
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            fn = lambda: o.gemm(a, w, bias=bias, out_op=out)
        elif kind == "res32":
            res = torch.randn(M, N, device="cuda")
            fn = lambda: o.gemm(a, w, bias=bias, residual=res, out_f32=res)
        else:
            out = torch.empty(M, N // 2, device="cuda", dtype=torch.bfloat16)
            fn = lambda: o.gemm(a, w, bias=bias, out_op=out, geglu=True)
        run(f"gemm {M}x{N}x{K} {kind} ({2.0 * M * N * K / 1e9:.1f} GFLOP)", fn)
    sys.exit(0)
# the level-0 (M = 131072 rows) GEMMs of the step: HBM-bound by the fp32 residual stream
for (M, N, K, kind) in [(131072, 320, 320, "op"), (131072, 320, 320, "res32"), (131072, 960, 320, "op"),
                        (131072, 320, 1280, "res32"), (131072, 2560, 320, "geglu"), (32768, 640, 640, "res32"),
                        (32768, 5120, 640, "geglu")]:
    a, w = bf(M, K), bf(N, K, scale=K ** -0.5)
    bias = torch.randn(N, device="cuda")
    if kind == "op":
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fn = lambda: o.gemm(a, w, bias=bias, out_op=out)
        nbytes = a.numel() * 2 + out.numel() * 2
    elif kind == "res32":
        res = torch.randn(M, N, device="cuda")
        fn = lambda: o.gemm(a, w, bias=bias, residual=res, out_f32=res)
        nbytes = a.numel() * 2 + res.numel() * 8
    else:
        out = torch.empty(M, N // 2, device="cuda", dtype=torch.bfloat16)
        fn = lambda: o.gemm(a, w, bias=bias, out_op=out, geglu=True)
        nbytes = a.numel() * 2 + out.numel() * 2
    run(f"gemm {M}x{N}x{K} {kind} ({nbytes / 1e6:.0f} MB, {2.0 * M * N * K / 1e9:.0f} GFLOP)", fn)

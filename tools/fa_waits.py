#!/usr/bin/env python
"""Where a softmax warp of the ping-pong attention kernel spends its cycles (counters written by the kernel itself;
needs the -DB200AD_GEMM_PROFILE build):   B200AD_LIB=.../libb200ad_prof.so python tools/fa_waits.py"""
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
for (Ln, d, dp) in [(4096, 40, 64), (1024, 80, 128)]:
    hd = 8 * dp
    qkv = (torch.randn(32 * Ln, 3 * hd, device="cuda")).to(torch.bfloat16)
    qkv.view(32 * Ln, 3, 8, dp)[:, 2, :, d] = 1.0
    for vs in (False, True):
        fn = lambda: o.attention(qkv[:, :hd], qkv[:, hd:2 * hd], qkv[:, 2 * hd:], 32, 8, Ln, Ln, d, dp, 1, v_ones=vs)
        fn(); fn()
        L.b200ad_debug_buffer(dbg.data_ptr()); dbg.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        L.b200ad_debug_buffer(None)
        c = dbg[:16].cpu().tolist()
        for name, v in (("tile A", c[:8]), ("tile B", c[8:])):
            n = max(v[7], 1)
            print(f"L={Ln} d={d} v_ones={vs} {e0.elapsed_time(e1)*1e3:.0f} us | {name}: {v[0]/n:.0f} cycles per key tile = "
                  f"wait S {v[1]/n:.0f} + tcgen05.ld {v[2]/n:.0f} + max {v[3]/n:.0f} + exp {v[4]/n:.0f} + wait PV(j-1) {v[5]/n:.0f} "
                  f"+ rescale/pack/store {v[6]/n:.0f}", flush=True)
        tl = dbg[64:64 + 48].cpu().tolist()
        t0 = min(x for x in tl if x > 0) if any(x > 0 for x in tl) else 0
        for t in range(3):
            row = tl[t * 16:(t + 1) * 16]
            if any(row):
                print(f"    exp phases of key tiles 8..15, query tile {'ABC'[t]} (cycles from the first): " +
                      " ".join(f"[{row[2*i]-t0}-{row[2*i+1]-t0}]" for i in range(8)), flush=True)

#!/usr/bin/env python
"""Algorithmic FLOPs of one UNetMotionModel forward (SURVEY.md Appendix C enumerator): 2*M*N*K per conv / Linear as
the reference computes it, 4*Bt*Lq*Lk*C per attention (unpadded head width and key count).
    python tools/flops.py            # the three reference sizes: 16f 256^2, 16f 512^2, 32f 768^2 at CFG batch 2
    from tools.flops import unet_flops; unet_flops(B=2, F=16, r=64)["total"]
"""
from collections import OrderedDict

L, X = 77, 768
CH = [320, 640, 1280, 1280]


def unet_flops(B=2, F=16, r=64):
    f = OrderedDict()

    def add(k, v):
        f[k] = f.get(k, 0.0) + v

    n = B * F

    def conv(cin, cout, r_out):
        add("conv3x3", 2.0 * n * r_out * r_out * 9 * cin * cout)

    def resnet(cin, cout, res):
        conv(cin, cout, res)
        conv(cout, cout, res)
        add("temb projections", 2.0 * n * 1280 * cout)
        if cin != cout:
            add("1x1 conv shortcuts", 2.0 * n * res * res * cin * cout)

    def xformer(c, res):
        hw = res * res
        add("spatial proj in/out (1x1)", 4.0 * n * hw * c * c)
        add("spatial self q,k,v,o", 8.0 * n * hw * c * c)
        add("spatial self-attention (QK^T+PV)", 4.0 * n * hw * hw * c)
        add("cross-attn q, o", 4.0 * n * hw * c * c)
        add("cross-attn k, v of text (per frame, redundant)", 4.0 * n * L * X * c)
        add("cross-attention (QK^T+PV, Lk=77)", 4.0 * n * hw * L * c)
        add("spatial FF", 24.0 * n * hw * c * c)

    def motion(c, res):
        hw = res * res
        add("temporal proj in/out", 4.0 * n * hw * c * c)
        add("temporal q,k,v,o (x2 attn)", 16.0 * n * hw * c * c)
        add("temporal attention (QK^T+PV)", 8.0 * (B * hw) * F * F * c)
        add("temporal FF (GEGLU up + down)", 24.0 * n * hw * c * c)

    res = r
    conv(4, CH[0], res)
    skips = [CH[0]]
    prev = CH[0]
    for i, c in enumerate(CH):
        for _ in range(2):
            resnet(prev, c, res)
            if i < 3:
                xformer(c, res)
            motion(c, res)
            prev = c
            skips.append(c)
        if i < 3:
            res //= 2
            conv(c, c, res)
            skips.append(c)
    resnet(prev, prev, res)
    xformer(prev, res)
    motion(prev, res)
    resnet(prev, prev, res)
    for i, c in enumerate([1280, 1280, 640, 320]):
        for _ in range(3):
            resnet(prev + skips.pop(), c, res)
            if i > 0:
                xformer(c, res)
            motion(c, res)
            prev = c
        if i < 3:
            res *= 2
            conv(c, c, res)
    conv(CH[0], 4, res)
    f["total"] = sum(f.values())
    return f


if __name__ == "__main__":
    for name, (B, F, r) in {"C1 16f 256^2": (2, 16, 32), "C2 16f 512^2": (2, 16, 64), "C5 32f 768^2": (2, 32, 96)}.items():
        fl = unet_flops(B, F, r)
        print(f"{name}: {fl['total'] / 1e9:.1f} GFLOP per CFG step")
        for k, v in sorted(fl.items(), key=lambda kv: -kv[1]):
            if k != "total":
                print(f"    {k:48s} {v / 1e9:10.1f}  {100 * v / fl['total']:5.1f} %")

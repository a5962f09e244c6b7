#!/usr/bin/env python
"""Run-to-run determinism stress of the fused UNet forward (full architecture, random init on the GPU): the same
inputs are evaluated ROUNDS times per kernel mode, interleaved with forwards of other shapes (to vary what the caching
allocator hands out), and every layer output is compared BITWISE with the first evaluation of that mode.  The first
layer that differs -- and which frame-images differ -- names the kernel whose result depends on anything but its inputs.

    python tools/stress_determinism.py [--latent 64] [--rounds 12]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import b200ad  # noqa: E402,F401
from b200ad.unet import UNetMotionConfig, UNetMotionModel  # noqa: E402

KINDS = ("ResnetBlock2D", "Transformer2DModel", "AnimateDiffTransformer3D", "Downsample2D", "Upsample2D")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--latent", type=int, default=64)
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--rounds", type=int, default=12)
    a = ap.parse_args()
    torch.manual_seed(0)
    with torch.device("cuda"):
        m = UNetMotionModel(UNetMotionConfig())
    g = torch.Generator().manual_seed(42)
    lat = torch.randn(1, 4, a.frames, a.latent, a.latent, generator=g).expand(2, -1, -1, -1, -1).contiguous().cuda()
    txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43)).cuda()
    lat_s = torch.randn(2, 4, a.frames, 32, 32, generator=g).cuda()
    cur = {}

    def wrap(name, mod):
        orig = mod.run

        def run(*args, **kw):
            out = orig(*args, **kw)
            x = out[0] if isinstance(out, tuple) else out
            cur[name] = x.clone()
            return out
        mod.run = run
    for n, mod in m.named_modules():
        if type(mod).__name__ in KINDS:
            wrap(n, mod)
    ref = {}
    bad = 0
    for r in range(a.rounds):
        for mode in ("bf16", "fp32"):
            m.set_mode(mode)
            cur.clear()
            out = m(lat, 961, encoder_hidden_states=txt).sample
            torch.cuda.synchronize()
            cur["__eps__"] = out.clone()
            if mode not in ref:
                ref[mode] = dict(cur)
                print(f"round {r} {mode}: reference kept ({len(cur)} tensors)", flush=True)
                continue
            first = None
            for name, t in cur.items():
                if not torch.equal(t, ref[mode][name]):
                    first = name
                    break
            if first is None:
                print(f"round {r} {mode}: bitwise identical", flush=True)
            else:
                bad += 1
                t, q = cur[first], ref[mode][first]
                nimg = 2 * a.frames
                d = (t.float() - q.float()).reshape(nimg, -1) if first != "__eps__" else (t - q).permute(0, 2, 1, 3, 4).reshape(nimg, -1)
                per = d.abs().amax(1)
                rel = float(d.norm() / q.float().norm())
                print(f"round {r} {mode}: FIRST DIFFERENCE at {first}: rel-L2 {rel:.3e}; images that differ: "
                      f"{[i for i in range(nimg) if per[i] > 0]}", flush=True)
                e = (cur["__eps__"] - ref[mode]["__eps__"])
                print(f"         final eps per half vs reference: "
                      f"{[float(e[i].norm() / ref[mode]['__eps__'][i].norm()) for i in range(2)]}", flush=True)
            # vary the allocator state between evaluations: other shapes, in both modes
            m.set_mode("bf16" if r % 2 else "fp32")
            keep = dict(cur)
            m(lat_s, 925, encoder_hidden_states=txt)
            cur.clear()
            del keep
    print(f"{bad} differing evaluations out of {2 * a.rounds - 2}")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())

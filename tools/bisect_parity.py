#!/usr/bin/env python
"""Layer-by-layer parity bisect on the GPU box: product UNet (fused .run() path) vs the CPU oracle, per CFG half.

    python tools/bisect_parity.py [--latent 64] [--frames 16] [--mode fp32|bf16] [--t 961]

The oracle runs once with forward hooks that keep the output of every resnet / spatial transformer / motion module /
down- and up-sampler; the product's ``run`` methods are wrapped to compare their row-matrix outputs against those
(relative L2 per CFG half).  The first line whose error jumps names the broken kernel's caller.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import b200ad  # noqa: E402,F401
from b200ad import unet as U  # noqa: E402
from oracle.unet_motion import make_oracle_unet  # noqa: E402


KINDS = ("ResnetBlock2D", "Transformer2DModel", "AnimateDiffTransformer3D", "Downsample2D", "Upsample2D")


def layer_errors(prod, oracle, lat, txt, t, log=print):
    """Run the oracle once with forward hooks (outputs of every resnet / transformer / motion module / sampler kept),
    then the product's fused plan with its ``run`` methods wrapped: prints relative L2 per CFG half for every layer
    and returns [(name, err_half0, err_half1)].  Leaves `prod` unwrapped."""
    saved, handles = {}, []

    def keep(name):
        def hook(mod, inp, out):
            saved[name] = (out if torch.is_tensor(out) else out[0]).detach()
        return hook
    for n, m in oracle.named_modules():
        if type(m).__name__ in KINDS:
            handles.append(m.register_forward_hook(keep(n)))
    with torch.no_grad():
        ref = oracle(lat, t, encoder_hidden_states=txt).sample
    for h in handles:
        h.remove()

    def rel(a, b):
        return float((a - b).norm() / b.norm().clamp_min(1e-30))
    rows, wrapped = [], []

    def wrap(name, mod):
        orig = mod.run
        wrapped.append(mod)

        def run(*a, **kw):
            out = orig(*a, **kw)
            x = out[0] if isinstance(out, tuple) else out
            r = saved[name].cuda()
            n_, c_, h_, w_ = r.shape
            p = x.view(n_, h_, w_, c_).permute(0, 3, 1, 2)
            half = max(n_ // 2, 1)
            e0, e1 = rel(p[:half], r[:half]), rel(p[half:], r[half:]) if n_ > 1 else 0.0
            per = ((p - r).flatten(1).norm(dim=1) / r.flatten(1).norm(dim=1)).tolist()
            rows.append((name, e0, e1))
            log(f"{name:58s} half0 {e0:.3e}  half1 {e1:.3e}  worst img {max(per):.3e} @ {per.index(max(per))}")
            return out
        mod.run = run
    for n, m in prod.named_modules():
        if type(m).__name__ in KINDS:
            wrap(n, m)
    try:
        out = prod(lat.cuda(), t, encoder_hidden_states=txt.cuda()).sample
    finally:
        for m in wrapped:
            del m.run
    log(f"final eps per half: {[rel(out[i].cpu(), ref[i]) for i in range(out.shape[0])]}")
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--latent", type=int, default=64)
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--mode", default="fp32")
    ap.add_argument("--t", type=float, default=961)
    ap.add_argument("--swap-text", action="store_true")
    ap.add_argument("--pre", action="store_true", help="first run config-1-shape forwards in bf16 and fp32 (test order)")
    ap.add_argument("--oracle-halves", action="store_true", help="also run the oracle on each half alone (B=1)")
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    oracle = make_oracle_unet(seed=0)
    prod = U.UNetMotionModel(U.UNetMotionConfig(), mode=args.mode)
    prod.load_state_dict(oracle.state_dict())
    prod.cuda()
    F, hw = args.frames, args.latent
    lat = torch.randn(1, 4, F, hw, hw, generator=torch.Generator().manual_seed(42)).expand(2, -1, -1, -1, -1).contiguous()
    txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43))
    if args.swap_text:
        txt = txt.flip(0).contiguous()

    if args.pre:
        l1 = torch.randn(1, 4, F, 32, 32, generator=torch.Generator().manual_seed(42)).expand(2, -1, -1, -1, -1).contiguous()
        for m in ("bf16", "fp32", "bf16", args.mode):
            prod.set_mode(m)
            prod(l1.cuda(), 925, encoder_hidden_states=txt.cuda())
        with torch.no_grad():
            oracle(l1, 925, encoder_hidden_states=txt)
    kinds = ("ResnetBlock2D", "Transformer2DModel", "AnimateDiffTransformer3D", "Downsample2D", "Upsample2D")
    saved = {}

    def keep(name):
        def hook(mod, inp, out):
            saved[name] = (out if torch.is_tensor(out) else out[0]).detach()
        return hook
    for n, m in oracle.named_modules():
        if type(m).__name__ in kinds:
            m.register_forward_hook(keep(n))
    with torch.no_grad():
        ref = oracle(lat, args.t, encoder_hidden_states=txt).sample
    print(f"oracle done: {len(saved)} module outputs kept", flush=True)
    if args.oracle_halves:
        keep_saved = dict(saved)
        with torch.no_grad():
            for i in range(2):
                r1 = oracle(lat[i:i + 1], args.t, encoder_hidden_states=txt[i:i + 1]).sample
                d = float((r1[0] - ref[i]).norm() / ref[i].norm())
                print(f"oracle: B=1 run of half {i} vs its own B=2 run: rel-L2 {d:.3e}", flush=True)
        saved.clear()
        saved.update(keep_saved)

    def rel(a, b):
        return float((a - b).norm() / b.norm().clamp_min(1e-30))

    order, wrapped = [], []

    def wrap(name, mod):
        orig = mod.run
        wrapped.append(mod)

        def run(*a, **kw):
            out = orig(*a, **kw)
            x = out[0] if isinstance(out, tuple) else out
            r = saved[name].cuda()                       # [N, C, h, w]
            n_, c_, h_, w_ = r.shape
            p = x.view(n_, h_, w_, c_).permute(0, 3, 1, 2)
            half = n_ // 2
            e0, e1 = rel(p[:half], r[:half]), rel(p[half:], r[half:])
            # worst image of each half
            per = ((p - r).flatten(1).norm(dim=1) / r.flatten(1).norm(dim=1)).tolist()
            order.append((name, e0, e1, max(per[:half]), max(per[half:])))
            print(f"{name:58s} half0 {e0:.3e}  half1 {e1:.3e}  worst img {max(per):.3e} @ {per.index(max(per))}", flush=True)
            return out
        mod.run = run
    for n, m in prod.named_modules():
        if type(m).__name__ in kinds:
            wrap(n, m)
    out = prod(lat.cuda(), args.t, encoder_hidden_states=txt.cuda()).sample
    print("final eps per half:", rel(out[0].cpu(), ref[0]), rel(out[1].cpu(), ref[1]))
    for m in wrapped:
        del m.run                                  # back to the class method
    # B = 1 runs of each half on their own
    for i in range(2):
        o1 = prod(lat[i:i + 1].cuda(), args.t, encoder_hidden_states=txt[i:i + 1].cuda()).sample
        print(f"B=1 run of half {i}: vs oracle {rel(o1[0].cpu(), ref[i]):.3e}; vs the B=2 product run {rel(o1[0], out[i]):.3e}")


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Writes tests/golden/oracle_small.json and tests/golden/schedulers.json from the CPU oracle (oracle/).

The reference's arithmetic for this path lives in diffusers, which is not installable in this image and holds no
stored tensors in /root/reference, so these fixtures are generated from the oracle restatement
itself: they pin it against accidental edits; the independent known-answers (param/module counts, DDIM tables,
the one-step KAT) are asserted directly in tests/test_oracle_cpu.py.
    python tests/golden/make_golden.py
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.schedulers import OracleDDIM, OracleEuler, denoise_loop  # noqa: E402
from oracle.unet_motion import UNetMotionConfig, make_oracle_unet  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
cfg = dict(block_out_channels=[32, 64, 64, 64], norm_num_groups=8, motion_norm_num_groups=8,
           num_attention_heads=2, motion_num_attention_heads=2, cross_attention_dim=48)
F, HW, T, G, STEPS, SEED = 4, 8, 925, 7.5, 3, 0
m = make_oracle_unet(seed=SEED, cfg=UNetMotionConfig(**{k: tuple(v) if isinstance(v, list) else v for k, v in cfg.items()}))
lat = torch.randn(1, 4, F, HW, HW, generator=torch.Generator().manual_seed(42))
txt = torch.randn(2, 77, 48, generator=torch.Generator().manual_seed(43))
with torch.no_grad():
    eps = m(torch.cat([lat] * 2), T, encoder_hidden_states=txt).sample
flat = eps.flatten()
idx = torch.linspace(0, flat.numel() - 1, 16).long()
sch = OracleDDIM()
final = denoise_loop(m, sch, lat, txt, G, STEPS)
json.dump({"seed": SEED, "cfg": cfg, "frames": F, "latent": HW, "timestep": T, "guidance": G, "steps": STEPS,
           "probe_index": idx.tolist(), "probe_value": flat[idx].tolist(),
           "sumsq": eps.double().pow(2).sum().item(), "final_sumsq": final.double().pow(2).sum().item(),
           "timesteps": sch.timesteps.tolist()}, open(os.path.join(HERE, "oracle_small.json"), "w"), indent=1)

out = {"ddim": {}, "euler": {}}
for n in (15, 20, 25, 30, 40, 50):
    d = OracleDDIM(); d.set_timesteps(n)
    ac = d.alphas_cumprod.double()
    cx, ce = [], []
    for t in d.timesteps.tolist():
        p = t - 1000 // n
        a_t, a_p = ac[t].item(), (ac[p].item() if p >= 0 else ac[0].item())
        cx.append((a_p / a_t) ** 0.5)
        ce.append((1 - a_p) ** 0.5 - (a_p * (1 - a_t) / a_t) ** 0.5)
    out["ddim"][str(n)] = {"timesteps": d.timesteps.tolist(), "cx": cx, "ce": ce}
    e = OracleEuler(); e.set_timesteps(n)
    out["euler"][str(n)] = {"timesteps": e.timesteps.tolist(), "sigmas": e.sigmas.tolist()}
json.dump(out, open(os.path.join(HERE, "schedulers.json"), "w"))
print("wrote golden fixtures")

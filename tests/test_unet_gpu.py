"""UNet-level parity (-m gpu): b200ad UNetMotionModel (sm_100a kernels through the C ABI) against the CPU oracle
(oracle/unet_motion.py) on identical bf16-representable random-init weights, latents and text embeddings.

Gates (BASELINE.json north_star): per-UNet-step noise-prediction rel-L2 <= 1e-2 in bf16, <= 1e-4 in the fp32
kernel mode.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = {"bf16": 1e-2, "fp32": 1e-4}


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


def _pair(cfg_kwargs, seed=0):
    from oracle.unet_motion import UNetMotionConfig as OC, make_oracle_unet
    from b200ad.unet import UNetMotionConfig as PC, UNetMotionModel
    oracle = make_oracle_unet(seed=seed, cfg=OC(**cfg_kwargs))
    prod = UNetMotionModel(PC(**cfg_kwargs))
    prod.load_state_dict(oracle.state_dict())
    return oracle, prod


def _inputs(B, F, h, w, seed=42):
    g = torch.Generator().manual_seed(seed)
    lat = torch.randn(1, 4, F, h, w, generator=g).expand(B, -1, -1, -1, -1).contiguous()
    txt = torch.randn(B, 77, 768, generator=torch.Generator().manual_seed(seed + 1))
    return lat, txt


SMALL = dict(block_out_channels=(64, 128, 256, 256))


@pytest.mark.parametrize("mode", ["bf16", "fp32"])
@pytest.mark.parametrize("B,F,h,w,t", [(2, 16, 16, 16, 925), (1, 8, 8, 24, 1), (2, 5, 16, 8, 500.5),
                                       # BASELINE configs[4] geometry in small: 32 frames (the PE-table limit), latent
                                       # sides 48/24/12/6 (768^2 gives 96/48/24/12: non-power-of-two token counts)
                                       (1, 32, 48, 48, 961)])
def test_small_unet_matches_oracle(mode, B, F, h, w, t):
    oracle, prod = _pair(SMALL)
    prod.set_mode(mode)
    lat, txt = _inputs(B, F, h, w)
    ref = oracle(lat, t, encoder_hidden_states=txt).sample
    out = prod(lat.cuda(), t, encoder_hidden_states=txt.cuda()).sample
    assert out.shape == ref.shape
    e = rel(out.cpu(), ref)
    print(f"small unet {mode} B={B} F={F} {h}x{w} t={t}: rel-L2 {e:.3e}")
    assert e < TOL[mode]


def test_small_unet_text_batch_conventions_and_timestep_forms():
    oracle, prod = _pair(SMALL)
    lat, txt = _inputs(2, 4, 8, 8)
    base = prod(lat.cuda(), 500, encoder_hidden_states=txt.cuda()).sample
    rep = prod(lat.cuda(), torch.tensor([500.0], dtype=torch.float16).cuda(),
               encoder_hidden_states=txt.repeat_interleave(4, 0).cuda()).sample
    assert torch.equal(base, rep)                      # [B*F,77,768] convention and fp16 [1] timestep: same bits
    rep2 = prod(lat.cuda(), torch.tensor(500).cuda(), encoder_hidden_states=txt.cuda()).sample
    assert torch.equal(base, rep2)
    prod.cache_text_kv(True)
    t = txt.cuda()
    a = prod(lat.cuda(), 500, encoder_hidden_states=t).sample
    b = prod(lat.cuda(), 500, encoder_hidden_states=t).sample      # second call served from the K/V cache
    assert torch.equal(a, base) and torch.equal(b, base)


def test_full_architecture_config1_step0():
    """BASELINE config 1 shape (16 frames, 256x256 -> 32x32 latents), CFG batch 2, t = 925 (step 0 of 15)."""
    oracle, prod = _pair({})
    lat, txt = _inputs(2, 16, 32, 32)
    torch.set_num_threads(max(1, torch.get_num_threads()))
    ref = oracle(lat, 925, encoder_hidden_states=txt).sample
    prod.cuda()
    out = prod(lat.cuda(), 925, encoder_hidden_states=txt.cuda()).sample
    e = rel(out.cpu(), ref)
    eh = [rel(out[i].cpu(), ref[i]) for i in range(2)]
    print(f"full unet bf16 config1 t=925: rel-L2 {e:.3e} (uncond {eh[0]:.3e}, cond {eh[1]:.3e})")
    assert e < TOL["bf16"]
    del oracle
    prod.set_mode("fp32")
    out32 = prod(lat.cuda(), 925, encoder_hidden_states=txt.cuda()).sample
    e32 = rel(out32.cpu(), ref)
    print(f"full unet fp32-kernel-mode config1 t=925: rel-L2 {e32:.3e}")
    assert e32 < TOL["fp32"]

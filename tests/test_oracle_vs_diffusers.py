"""Closes the "parity unpinned" gap the moment an image provides the real package: load the oracle's
state dict into diffusers' own UNetMotionModel / AutoencoderKL / schedulers and compare one forward in fp32.
Skipped when ``diffusers`` is not importable (this image: no wheel, no network) or when the importable ``diffusers`` is
this repo's own shim (compat/diffusers)."""
import os

import pytest
import torch

diffusers = pytest.importorskip("diffusers")
if "b200ad" in getattr(diffusers, "__version__", "") or "compat" in os.path.dirname(getattr(diffusers, "__file__", "")):
    pytest.skip("the importable `diffusers` is the b200ad shim, not the reference's dependency", allow_module_level=True)


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


def test_unet_motion_model_matches_oracle():
    from diffusers import MotionAdapter, UNet2DConditionModel, UNetMotionModel
    from oracle.unet_motion import make_oracle_unet
    oracle = make_oracle_unet(seed=0, round_to=None)
    base = UNet2DConditionModel(sample_size=64, in_channels=4, out_channels=4, block_out_channels=(320, 640, 1280, 1280),
                                layers_per_block=2, cross_attention_dim=768, attention_head_dim=8,
                                down_block_types=("CrossAttnDownBlock2D",) * 3 + ("DownBlock2D",),
                                up_block_types=("UpBlock2D",) + ("CrossAttnUpBlock2D",) * 3)
    adapter = MotionAdapter(block_out_channels=(320, 640, 1280, 1280), motion_layers_per_block=2,
                            motion_mid_block_layers_per_block=1, motion_num_attention_heads=8,
                            motion_norm_num_groups=32, motion_max_seq_length=32, use_motion_mid_block=True)
    ref = UNetMotionModel.from_unet2d(base, adapter).eval()
    missing, unexpected = ref.load_state_dict(oracle.state_dict(), strict=False)
    assert not unexpected and all("pos_embed" in k for k in missing), (missing[:5], unexpected[:5])
    lat = torch.randn(2, 4, 16, 32, 32, generator=torch.Generator().manual_seed(42))
    txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43))
    with torch.no_grad():
        a = ref(lat, 925, encoder_hidden_states=txt).sample
        b = oracle(lat, 925, encoder_hidden_states=txt).sample
    assert rel(b, a) < 1e-5


def test_schedulers_match_oracle():
    from diffusers import DDIMScheduler, EulerDiscreteScheduler
    from oracle.schedulers import OracleDDIM, OracleEuler
    cfg = dict(num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012, beta_schedule="linear")
    d = DDIMScheduler(**cfg, steps_offset=1, clip_sample=False, set_alpha_to_one=False)
    o = OracleDDIM()
    x, e = torch.randn(1, 4, 16, 8, 8), torch.randn(1, 4, 16, 8, 8)
    for n in (15, 25, 50):
        d.set_timesteps(n)
        o.set_timesteps(n)
        assert torch.equal(d.timesteps.cpu(), o.timesteps)
        t = int(o.timesteps[3])
        assert rel(o.step(e, t, x), d.step(e, t, x).prev_sample) < 1e-6
    eu = EulerDiscreteScheduler(**cfg, timestep_spacing="linspace")
    oe = OracleEuler()
    eu.set_timesteps(25)
    oe.set_timesteps(25)
    assert torch.allclose(eu.sigmas.cpu(), oe.sigmas) and abs(float(eu.init_noise_sigma) - oe.init_noise_sigma) < 1e-6


def test_vae_decoder_matches_oracle():
    from diffusers import AutoencoderKL
    from oracle.vae import make_oracle_vae
    oracle = make_oracle_vae(seed=1, round_to=None)
    ref = AutoencoderKL(in_channels=3, out_channels=3, latent_channels=4, block_out_channels=(128, 256, 512, 512),
                        down_block_types=("DownEncoderBlock2D",) * 4, up_block_types=("UpDecoderBlock2D",) * 4,
                        layers_per_block=2).eval()
    missing, unexpected = ref.load_state_dict(oracle.state_dict(), strict=False)
    assert not unexpected and all(k.startswith(("encoder.", "quant_conv.")) for k in missing)
    z = torch.randn(2, 4, 16, 16, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        assert rel(oracle.decode(z), ref.decode(z).sample) < 1e-5

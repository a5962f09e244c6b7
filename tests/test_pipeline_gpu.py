"""Pipeline-level parity (-m gpu): the CUDA-graph denoise loop of b200ad.AnimateDiffPipeline against the oracle's
restated AnimateDiffPipeline loop (oracle/schedulers.py denoise_loop) on a reduced-width UNet of the same
architecture, identical weights / latents / text.  DDIM (05_grid_search_ablation.py:136-141) and Euler
(01_baseline_generation.py:76-80) configurations."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SMALL = dict(block_out_channels=(64, 128, 256, 256))


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


def _make(mode="bf16"):
    from oracle.unet_motion import UNetMotionConfig as OC, make_oracle_unet
    from b200ad.unet import UNetMotionConfig as PC, UNetMotionModel
    from b200ad.pipeline import AnimateDiffPipeline
    from b200ad.schedulers import PNDMPlaceholder, SD15_SCHEDULER_CONFIG
    oracle = make_oracle_unet(seed=0, cfg=OC(**SMALL))
    unet = UNetMotionModel(PC(**SMALL), mode=mode)
    unet.load_state_dict(oracle.state_dict())
    pipe = AnimateDiffPipeline(unet.cuda(), PNDMPlaceholder(SD15_SCHEDULER_CONFIG), vae=None)
    return oracle, pipe


@pytest.mark.parametrize("sched", ["ddim", "euler"])
@pytest.mark.parametrize("graph", [True, False])
def test_denoise_loop_matches_oracle(sched, graph):
    from b200ad.schedulers import DDIMScheduler, EulerDiscreteScheduler
    from oracle.schedulers import OracleDDIM, OracleEuler, denoise_loop
    oracle, pipe = _make("fp32")               # fp32 kernel mode: the loop logic must agree to ~1e-5
    pipe.use_graph = graph
    if sched == "ddim":
        pipe.scheduler = DDIMScheduler.from_config(pipe.scheduler.config, beta_schedule="linear", steps_offset=1,
                                                   clip_sample=False)
        osch = OracleDDIM()
    else:
        pipe.scheduler = EulerDiscreteScheduler.from_config(pipe.scheduler.config, timestep_spacing="linspace",
                                                            beta_schedule="linear")
        osch = OracleEuler()
    n, g = 4, 7.5
    lat = torch.randn(1, 4, 8, 8, 8, generator=torch.Generator().manual_seed(42))
    txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43))
    osch.set_timesteps(n)
    ref = denoise_loop(oracle, osch, lat.clone(), txt, g, n)
    out = pipe(prompt_embeds=txt[1:], negative_prompt_embeds=txt[:1], num_frames=8, height=64, width=64,
               num_inference_steps=n, guidance_scale=g, latents=lat.clone(), output_type="latent").frames
    e = rel(out.cpu(), ref)
    print(f"{sched} graph={graph}: {n}-step trajectory rel-L2 {e:.3e}")
    assert e < 2e-4
    # a second video through the same captured graph, different prompt + latents + guidance
    lat2 = torch.randn(1, 4, 8, 8, 8, generator=torch.Generator().manual_seed(7))
    txt2 = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(8))
    osch.set_timesteps(n)
    ref2 = denoise_loop(oracle, osch, lat2.clone(), txt2, 5.0, n)
    out2 = pipe(prompt_embeds=txt2[1:], negative_prompt_embeds=txt2[:1], num_frames=8, height=64, width=64,
                num_inference_steps=n, guidance_scale=5.0, latents=lat2.clone(), output_type="latent").frames
    assert rel(out2.cpu(), ref2) < 2e-4


def test_bf16_trajectory_and_string_prompts():
    from b200ad.schedulers import DDIMScheduler
    from b200ad.pipeline import synthetic_text_embeds
    from oracle.schedulers import OracleDDIM, denoise_loop
    oracle, pipe = _make("bf16")
    pipe.scheduler = DDIMScheduler.from_config(pipe.scheduler.config, beta_schedule="linear", steps_offset=1,
                                               clip_sample=False)
    gen = torch.manual_seed(42)                 # the reference passes the global CPU generator (05:156)
    out = pipe(prompt="a cat", negative_prompt="blurry", num_frames=8, height=128, width=128,
               num_inference_steps=5, guidance_scale=7.5, generator=gen, output_type="latent").frames
    torch.manual_seed(42)
    lat = torch.randn(1, 4, 8, 16, 16)
    txt = torch.stack([synthetic_text_embeds("blurry"), synthetic_text_embeds("a cat")])
    osch = OracleDDIM()
    osch.set_timesteps(5)
    ref = denoise_loop(oracle, osch, lat, txt, 7.5, 5)
    e = rel(out.cpu(), ref)
    print(f"bf16 5-step DDIM trajectory rel-L2 {e:.3e}")
    assert e < 3e-2          # guided-eps error is CFG-amplified; per-UNet-call gate is 1e-2

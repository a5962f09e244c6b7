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


@pytest.mark.parametrize("mode", ["bf16", "fp32"])
@pytest.mark.parametrize("B,F,h,w", [(2, 16, 64, 64), (2, 8, 8, 8), (1, 5, 24, 40), (2, 32, 96, 96)])
def test_no_read_before_write(mode, B, F, h, w):
    """Every buffer the kernels' front end allocates is filled with NaN instead of being left uninitialised
    (ops.POISON): a kernel that reads bytes no kernel of this forward has written -- a statistics slot, a pad column, a
    tile tail -- turns the noise prediction into NaN instead of silently picking up what the caching allocator last
    kept there (which made a parity test depend on the tests that ran before it).  Shapes: the benchmarked token
    counts (64x64), several images per conv tile (8x8), ragged boxes (24x40), config 5's 96x96 x 32 frames."""
    import b200ad.ops as O
    from b200ad.unet import UNetMotionConfig as PC, UNetMotionModel
    from b200ad.vae import AutoencoderKLDecoder
    torch.manual_seed(0)
    prod = UNetMotionModel(PC(**SMALL), mode=mode).cuda()
    lat, txt = _inputs(B, F, h, w)
    clean = prod(lat.cuda(), 500, encoder_hidden_states=txt.cuda()).sample
    O.POISON = True
    try:
        out = prod(lat.cuda(), 500, encoder_hidden_states=txt.cuda()).sample
        assert torch.isfinite(out).all(), f"{int((~torch.isfinite(out)).sum())} non-finite outputs: a kernel read unwritten memory"
        assert torch.equal(out, clean)
        if (h, w) == (8, 8):
            vae = AutoencoderKLDecoder((128, 128, 64, 64), mode=mode).cuda()
            z = torch.randn(1, 4, 3, h, w, device="cuda")
            img_clean = None
            O.POISON = False
            img_clean = vae.decode_video(z)
            O.POISON = True
            img = vae.decode_video(z)
            assert torch.isfinite(img).all() and torch.equal(img, img_clean)
    finally:
        O.POISON = False


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


@pytest.fixture(scope="module")
def full_pair():
    """Full SD1.5 + motion-adapter architecture: the oracle (CPU, fp32) and the product with the oracle's
    bf16-representable weights.  Built once for all full-architecture tests (construction is ~1.3 G parameters)."""
    oracle, prod = _pair({})
    prod.cuda()
    return oracle, prod


def _gate(prod, oracle, lat, txt, t, what, modes=("bf16",)):
    ref = oracle(lat, t, encoder_hidden_states=txt).sample
    errs, outs = {}, {}
    for mode in modes:
        prod.set_mode(mode)
        out = prod(lat.cuda(), t, encoder_hidden_states=txt.cuda()).sample
        assert out.shape == ref.shape and torch.isfinite(out).all()
        errs[mode] = rel(out.cpu(), ref)
        outs[mode] = out
        halves = [rel(out[i].cpu(), ref[i]) for i in range(ref.shape[0])]
        print(f"full unet {mode} {what} t={t}: rel-L2 {errs[mode]:.3e} (per CFG half {['%.3e' % h for h in halves]})")
    for mode in modes:
        if not errs[mode] < TOL[mode]:
            # diagnostic: which layer's output leaves the oracle first (per CFG half), in THIS process state
            from tools.bisect_parity import layer_errors
            prod.set_mode(mode)
            again = prod(lat.cuda(), t, encoder_hidden_states=txt.cuda()).sample
            print(f"--- mode {mode}: a second forward differs from the first by rel-L2 {rel(again, outs[mode]):.3e}; "
                  f"its error vs the oracle is {rel(again.cpu(), ref):.3e}")
            print(f"--- layer bisect, mode {mode} ---")
            layer_errors(prod, oracle, lat, txt, t)
    prod.set_mode("bf16")
    for mode in modes:
        assert errs[mode] < TOL[mode], (what, mode, errs[mode])
    return ref


@pytest.mark.parametrize("t", [925, 463, 1])
def test_full_architecture_config1_three_timesteps(full_pair, t):
    """BASELINE config 1 shape (16 frames, 256x256 -> 32x32 latents), CFG batch 2, at the first, a middle and the last
    timestep of the 15-step DDIM schedule (SURVEY.md Appendix E.2 asks the gate at every step; three span the range).
    The fp32 kernel mode is gated at the first one."""
    oracle, prod = full_pair
    lat, txt = _inputs(2, 16, 32, 32)
    _gate(prod, oracle, lat, txt, t, "config1 (16f 256^2)", modes=("bf16", "fp32") if t == 925 else ("bf16",))


def test_full_architecture_config2_benchmarked_shape(full_pair):
    """BASELINE config 2 -- the shape bench.py times: 16 frames 512x512 -> latents [2,4,16,64,64], t = 961 (step 0 of
    25).  Level-0 self-attention is L = 4096, d = 40 over 32 images (the attn_fa three-tile instance), the GEMM / conv
    tile scheduler runs at M = 131072.  bf16 <= 1e-2 and fp32 kernel mode <= 1e-4 against the fp32 CPU oracle."""
    oracle, prod = full_pair
    lat, txt = _inputs(2, 16, 64, 64)
    _gate(prod, oracle, lat, txt, 961, "config2 (16f 512^2)", modes=("bf16", "fp32"))


@pytest.mark.parametrize("h", [32, 64])
def test_full_architecture_no_read_before_write(full_pair, h):
    """ops.POISON (see test_no_read_before_write) on the full architecture at the config-1 / config-2 shapes: the real
    head widths (40 / 80 / 160 -> attn_fa, attn_x, attn_tc), C = 320 .. 2560 GroupNorm / LayerNorm variants."""
    import b200ad.ops as O
    _, prod = full_pair
    lat, txt = _inputs(2, 16, h, h)
    for mode in ("bf16", "fp32"):
        prod.set_mode(mode)
        clean = prod(lat.cuda(), 961, encoder_hidden_states=txt.cuda()).sample
        O.POISON = True
        try:
            out = prod(lat.cuda(), 961, encoder_hidden_states=txt.cuda()).sample
        finally:
            O.POISON = False
        bad = int((~torch.isfinite(out)).sum())
        assert bad == 0, f"{mode} {h}x{h}: {bad} non-finite outputs -- a kernel read unwritten memory"
        assert torch.equal(out, clean), f"{mode} {h}x{h}: result depends on the previous content of its buffers"
    prod.set_mode("bf16")


def test_full_architecture_euler_bf16_one_step(full_pair):
    """EulerDiscreteScheduler (01_baseline_generation.py:76-80) in bf16 on the full architecture: one step of the
    25-step schedule through the CUDA-graph DenoisePlan (scale_model_input -> UNet x2 -> CFG -> Euler update) against the
    oracle's loop body.  Config-1 shape."""
    from b200ad.pipeline import AnimateDiffPipeline
    from b200ad.schedulers import EulerDiscreteScheduler, SD15_SCHEDULER_CONFIG
    from oracle.schedulers import OracleEuler, cfg_combine
    oracle, prod = full_pair
    g = 7.5
    lat = torch.randn(1, 4, 16, 32, 32, generator=torch.Generator().manual_seed(42))
    txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43))
    osch = OracleEuler()
    osch.set_timesteps(25)
    x = lat * osch.init_noise_sigma
    t = osch.timesteps[0]
    x_in = osch.scale_model_input(torch.cat([x] * 2), t)
    eps = oracle(x_in, t, encoder_hidden_states=txt).sample
    ref = osch.step(cfg_combine(eps, g), t, x)
    e_eps = rel(prod(x_in.cuda(), float(t), encoder_hidden_states=txt.cuda()).sample.cpu(), eps)
    print(f"full unet bf16 Euler-scaled input t={float(t):.1f}: noise-prediction rel-L2 {e_eps:.3e}")
    assert e_eps < TOL["bf16"]
    sch = EulerDiscreteScheduler.from_config(SD15_SCHEDULER_CONFIG, timestep_spacing="linspace", beta_schedule="linear")
    pipe = AnimateDiffPipeline(prod, sch, vae=None)
    sch.set_timesteps(25)
    plan = pipe._plan(2, 16, 32, 32)
    plan.load((lat * sch.init_noise_sigma).cuda(), txt.cuda(), sch, g)
    plan.capture()
    plan.step(0)
    torch.cuda.synchronize()
    # the latents move by dt*guided eps; compare the UPDATE (x' - x), which is what the UNet produced
    d_ref = ref - x
    d_out = plan.latents.cpu() - x
    e_upd, e_lat = rel(d_out, d_ref), rel(plan.latents.cpu(), ref)
    print(f"full unet bf16 Euler step 0 of 25: guided update rel-L2 {e_upd:.3e}, latents rel-L2 {e_lat:.3e}")
    assert e_lat < 1e-2
    # the guided eps is CFG-amplified, u + 7.5 (c - u): two halves at the 1e-2 per-call gate give up to
    # sqrt(6.5^2 + 7.5^2) * 1e-2 ~ 0.1 relative to |eps|; measured 4e-2
    assert e_upd < 0.1


def test_attention_processor_and_module_mode_motion_module(full_pair):
    """a8 / a9 surface: B200AttnProcessor on a product Attention vs the oracle Attention (self- and cross-attention),
    and ``unet.down_blocks[0].motion_modules[0](x, num_frames=16)`` (03_trace_forward_pass.py:182-205) vs the oracle
    motion module."""
    from b200ad.attention_processor import B200AttnProcessor
    oracle, prod = full_pair
    blk_o = oracle.down_blocks[1].attentions[0].transformer_blocks[0]
    blk_p = prod.down_blocks[1].attentions[0].transformer_blocks[0]
    x = torch.randn(3, 256, 640, generator=torch.Generator().manual_seed(5))
    ctx = torch.randn(3, 77, 768, generator=torch.Generator().manual_seed(6))
    proc = B200AttnProcessor()
    out = proc(blk_p.attn1, x.cuda())
    assert rel(out.cpu(), blk_o.attn1(x)) < 1e-2
    out = proc(blk_p.attn2, x.cuda(), encoder_hidden_states=ctx.cuda())
    assert rel(out.cpu(), blk_o.attn2(x, ctx)) < 1e-2
    # through the module's own forward with the processor installed via the UNet surface
    prod.set_attn_processor(B200AttnProcessor())
    assert len(prod.attn_processors) == 74
    out = blk_p.attn2(x.cuda(), encoder_hidden_states=ctx.cuda())
    assert rel(out.cpu(), blk_o.attn2(x, ctx)) < 1e-2
    mm_o, mm_p = oracle.down_blocks[0].motion_modules[0], prod.down_blocks[0].motion_modules[0]
    h = torch.randn(16, 320, 16, 16, generator=torch.Generator().manual_seed(7))
    out = mm_p(h.cuda(), num_frames=16)
    assert out.shape == h.shape
    e = rel(out.cpu() - h, mm_o(h, num_frames=16) - h)
    print(f"motion module (module mode) residual-branch rel-L2 {e:.3e}")
    assert e < 1e-2
    res_o, res_p = oracle.down_blocks[1].resnets[0], prod.down_blocks[1].resnets[0]
    h = torch.randn(4, 320, 16, 16, generator=torch.Generator().manual_seed(8))
    temb = torch.randn(4, 1280, generator=torch.Generator().manual_seed(9))
    assert rel(res_p(h.cuda(), temb.cuda()).cpu(), res_o(h, temb)) < 1e-2


def test_trajectory_gates_and_psnr_config1(full_pair):
    """The north star's three gates on BASELINE config 1 (16 frames 256x256, 15 DDIM steps, CFG 7.5), full
    architecture, bf16 kernels: (1) per-UNet-step noise-prediction rel-L2 <= 1e-2 at EVERY step of the oracle's
    trajectory, (2) DDIM timestep indexing bit-exact, (3) decoded frames PSNR >= 35 dB
    (the reference's formula, /root/reference/experiments/06_measure_grid_search.py:209-218) between the oracle's
    video (CPU loop + oracle VAE) and the product's (CUDA-graph loop + b200ad VAE), same VAE weights."""
    from b200ad.pipeline import AnimateDiffPipeline
    from b200ad.schedulers import DDIMScheduler, SD15_SCHEDULER_CONFIG
    from b200ad.vae import AutoencoderKLDecoder
    from oracle.schedulers import OracleDDIM, cfg_combine
    from oracle.vae import decode_latents, make_oracle_vae, postprocess_video, psnr
    oracle, prod = full_pair
    n, g = 15, 7.5
    lat0 = torch.randn(1, 4, 16, 32, 32, generator=torch.Generator().manual_seed(42))
    txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43))
    osch = OracleDDIM()
    osch.set_timesteps(n)
    sch = DDIMScheduler.from_config(SD15_SCHEDULER_CONFIG, beta_schedule="linear", steps_offset=1, clip_sample=False)
    sch.set_timesteps(n)
    assert sch.timesteps.dtype == torch.int64 and torch.equal(sch.timesteps, osch.timesteps)      # gate (2)
    x, worst = lat0.clone(), 0.0
    txt_d = txt.cuda()
    for i, t in enumerate(osch.timesteps):
        x_in = torch.cat([x] * 2)
        eps_o = oracle(x_in, t, encoder_hidden_states=txt).sample
        eps_p = prod(x_in.cuda(), int(t), encoder_hidden_states=txt_d).sample
        e = rel(eps_p.cpu(), eps_o)
        worst = max(worst, e)
        print(f"step {i:2d} t={int(t):3d}: noise-prediction rel-L2 {e:.3e}")
        assert e < TOL["bf16"], (i, e)                                                            # gate (1)
        x = osch.step(cfg_combine(eps_o, g), t, x)
    vae_o = make_oracle_vae(seed=1)
    vae_p = AutoencoderKLDecoder()
    vae_p.load_state_dict(vae_o.state_dict())
    pipe = AnimateDiffPipeline(prod, sch, vae=vae_p.cuda())
    res = pipe(prompt_embeds=txt[1:], negative_prompt_embeds=txt[:1], num_frames=16, height=256, width=256,
               num_inference_steps=n, guidance_scale=g, latents=lat0.clone(), output_type="pil")
    frames = res.frames[0]
    assert len(frames) == 16 and frames[0].size == (256, 256) and frames[0].mode == "RGB"
    import numpy as np
    vid_p = torch.from_numpy(np.stack([np.asarray(f) for f in frames])).float() / 255.0         # [16,256,256,3]
    vid_o = postprocess_video(decode_latents(vae_o, x)).float() / 255.0
    e_lat = rel(pipe.last_latents.cpu(), x)
    per_frame = [psnr(vid_p[f], vid_o[f]) for f in range(16)]
    print(f"15-step trajectory: worst per-step rel-L2 {worst:.3e}; final latents rel-L2 {e_lat:.3e}; "
          f"PSNR of decoded frames min {min(per_frame):.1f} dB / mean {sum(per_frame) / 16:.1f} dB")
    assert min(per_frame) >= 35.0                                                                  # gate (3)

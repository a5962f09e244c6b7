"""CPU tests (-m "not gpu"): the oracle against everything the reference pins for this path.

The reference has no tests and no stored tensors (parity unpinned); what it does publish are
structural known-answers (/root/reference/docs/02_video_diffusion_architecture.md:86-91, computed by
/root/reference/experiments/02_architecture_inspection.py:38-60) and the scheduler formulas
(/root/reference/docs/01_diffusion_fundamentals.md:109-136,182-186).  The golden numbers below are the ones
recorded in SURVEY.md Appendix B (computed from those formulas with torch CPU fp32) and tests/golden/*.json
(produced by tests/golden/make_golden.py from the oracle; they pin the oracle against accidental edits)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle.schedulers import OracleDDIM, OracleEuler, cfg_combine, denoise_loop
from oracle.unet_motion import UNetMotionConfig, UNetMotionModel, make_oracle_unet

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

DDIM_TABLE = {   # n -> (first four, last three, prev index at the last step)
    15: ([925, 859, 793, 727], [133, 67, 1], -65),
    20: ([951, 901, 851, 801], [101, 51, 1], -49),
    25: ([961, 921, 881, 841], [81, 41, 1], -39),
    30: ([958, 925, 892, 859], [67, 34, 1], -32),
    40: ([976, 951, 926, 901], [51, 26, 1], -24),
    50: ([981, 961, 941, 921], [41, 21, 1], -19),
}


@pytest.mark.parametrize("n", sorted(DDIM_TABLE))
def test_ddim_timesteps_bit_exact(n):
    first, last, prev = DDIM_TABLE[n]
    s = OracleDDIM()
    s.set_timesteps(n)
    ts = s.timesteps.tolist()
    assert s.timesteps.dtype == torch.int64 and len(ts) == n
    assert ts[:4] == first and ts[-3:] == last
    assert ts[-1] - 1000 // n == prev


def test_ddim_alphas_and_one_step_kat():
    s = OracleDDIM()
    ac = s.alphas_cumprod
    assert abs(ac[0].item() - 0.999149978) < 1e-7
    assert abs(ac[1].item() - 0.998289526) < 1e-7
    assert abs(ac[961].item() - 2.478332259e-03) < 1e-9
    assert abs(ac[999].item() - 1.578962780e-03) < 1e-9      # 4.66e-3 would mean scaled_linear was used
    s.set_timesteps(25)
    x = torch.tensor([1.5409960747, -0.2934288979, -2.1787893772, 0.5684312582])
    e = torch.tensor([-1.0845223665, -1.3985954523, 0.4033468366, 0.8380263448])
    out = s.step(e, 961, x)
    ref = torch.tensor([2.2153205872, -0.0092060566, -2.8417060375, 0.4988986850])
    assert torch.allclose(out, ref, rtol=0, atol=2e-6)


def test_ddim_last_step_uses_final_alpha():
    s = OracleDDIM()
    s.set_timesteps(15)
    x, e = torch.ones(3), torch.full((3,), 0.5)
    a_t, a_p = s.alphas_cumprod[1], s.alphas_cumprod[0]
    ref = a_p.sqrt() * (x - (1 - a_t).sqrt() * e) / a_t.sqrt() + (1 - a_p).sqrt() * e
    assert torch.allclose(s.step(e, 1, x), ref)


def test_euler_golden_n25():
    s = OracleEuler()
    s.set_timesteps(25)
    ts, sg = s.timesteps.numpy(), s.sigmas.numpy()
    assert np.allclose(ts[:3], [999, 957.375, 915.75]) and np.allclose(ts[-2:], [41.625, 0])
    assert np.allclose(sg[:3], [25.146116, 19.642946, 15.490523], rtol=1e-6)
    assert np.allclose(sg[-3:], [0.21734539, 0.02916753, 0.0], rtol=1e-6)
    assert abs(s.init_noise_sigma - 25.146116) < 1e-5
    x, e = torch.randn(8), torch.randn(8)
    assert torch.allclose(s.scale_model_input(x, ts[0]), x / (sg[0] ** 2 + 1) ** 0.5)
    out = s.step(e, ts[0], x)
    assert torch.allclose(out, x + (sg[1] - sg[0]) * e, atol=1e-5)
    assert s.step_index == 1


def test_cfg_combine_formula_and_row_order():
    eps = torch.stack([torch.full((4,), 1.0), torch.full((4,), 3.0)])     # row 0 = uncond, row 1 = cond
    assert torch.equal(cfg_combine(eps, 7.5), torch.full((1, 4), 1.0 + 7.5 * 2.0))


def test_structural_known_answers_full_architecture():
    """1312.7 M parameters, 639 temporal/motion-named modules, 32 spatial Attention modules
    (docs/02_video_diffusion_architecture.md:86-91) with the reference's exact predicates
    (experiments/02_architecture_inspection.py:38-60)."""
    with torch.device("meta"):
        m = UNetMotionModel(UNetMotionConfig())
    total = sum(p.numel() for p in m.parameters())
    assert total == 1_312_730_244
    motion = sum(p.numel() for n, p in m.named_parameters() if "motion_modules" in n)
    assert motion == 453_209_280 and total - motion == 859_520_964
    temporal, spatial = [], []
    for name, mod in m.named_modules():
        low = name.lower()
        if "temporal" in low or "motion" in low:
            temporal.append(name)
        elif "attn" in low and type(mod).__name__ == "Attention":
            spatial.append(name)
    assert len(temporal) == 639
    assert len(spatial) == 32
    a = m.down_blocks[0].attentions[0].transformer_blocks[0].attn1
    assert a.heads == 8 and a.to_q.in_features == 320 and a.to_q.out_features == 320
    assert type(m.down_blocks[0].motion_modules[0]).__name__ == "AnimateDiffTransformer3D"
    keys = set(m.state_dict().keys())
    for k in ("down_blocks.0.motion_modules.0.temporal_transformer.transformer_blocks.0.attn1.to_q.weight",
              "down_blocks.0.motion_modules.0.transformer_blocks.0.attn1.to_q.weight"):
        if k in keys:
            break
    else:
        raise AssertionError("motion-module state-dict keys do not follow the diffusers naming")
    assert "mid_block.attentions.0.transformer_blocks.0.attn2.to_k.weight" in keys
    assert "up_blocks.3.resnets.2.conv_shortcut.weight" in keys
    assert "time_embedding.linear_1.weight" in keys


SMALL = dict(block_out_channels=(32, 64, 64, 64), norm_num_groups=8, motion_norm_num_groups=8,
             num_attention_heads=2, motion_num_attention_heads=2, cross_attention_dim=48)


def _small_inputs(F=4, hw=8):
    lat = torch.randn(1, 4, F, hw, hw, generator=torch.Generator().manual_seed(42))
    txt = torch.randn(2, 77, 48, generator=torch.Generator().manual_seed(43))
    return lat, txt


def test_oracle_shapes_text_conventions_and_frame_coupling():
    m = make_oracle_unet(seed=0, cfg=UNetMotionConfig(**SMALL))
    lat, txt = _small_inputs()
    x2 = torch.cat([lat] * 2)
    with torch.no_grad():
        a = m(x2, 925, encoder_hidden_states=txt).sample
        b = m(x2, torch.tensor(925), encoder_hidden_states=txt.repeat_interleave(4, 0)).sample
        assert a.shape == x2.shape and torch.isfinite(a).all()
        assert torch.allclose(a, b, atol=1e-6)                 # [B,77,X] and [B*F,77,X] conventions agree
        assert (a[0] - a[1]).norm() / a[0].norm() > 1e-3       # the text path is live (halves differ)
        x3 = x2.clone()
        x3[:, :, 0] += 1.0                                     # perturb frame 0 only
        c = m(x3, 925, encoder_hidden_states=txt).sample
        assert (c[:, :, 3] - a[:, :, 3]).abs().max() > 1e-6    # motion modules couple the frames
        # CFG halves are independent through the whole UNet (the 2-GPU split is exact)
        h0 = m(lat, 925, encoder_hidden_states=txt[:1]).sample
        assert torch.allclose(h0[0], a[0], atol=2e-5)


def test_oracle_golden_fixture():
    """tests/golden/oracle_small.json was written by tests/golden/make_golden.py; a change in the oracle's
    arithmetic shows up here before it can silently move the GPU parity target."""
    with open(os.path.join(GOLDEN, "oracle_small.json")) as f:
        g = json.load(f)
    m = make_oracle_unet(seed=g["seed"], cfg=UNetMotionConfig(**{k: tuple(v) if isinstance(v, list) else v
                                                                for k, v in g["cfg"].items()}))
    lat, txt = _small_inputs(g["frames"], g["latent"])
    with torch.no_grad():
        eps = m(torch.cat([lat] * 2), g["timestep"], encoder_hidden_states=txt).sample
    flat = eps.flatten()
    idx = torch.tensor(g["probe_index"])
    assert torch.allclose(flat[idx], torch.tensor(g["probe_value"]), rtol=2e-4, atol=2e-5)
    assert abs(eps.double().pow(2).sum().item() - g["sumsq"]) / g["sumsq"] < 1e-4
    sch = OracleDDIM()
    final = denoise_loop(m, sch, lat, txt, g["guidance"], g["steps"])
    assert abs(final.double().pow(2).sum().item() - g["final_sumsq"]) / g["final_sumsq"] < 5e-4
    assert sch.timesteps.tolist() == g["timesteps"]


def test_oracle_clip_text_matches_transformers():
    """Row f2's oracle is PINNED: transformers (the reference's own dependency for the text encoder,
    /root/reference/requirements.txt:7) is importable here, so the restatement is compared with the library's
    CLIPTextModel on identical random-init weights and token ids."""
    transformers = pytest.importorskip("transformers")
    from oracle.clip_text import ClipTextConfig, make_oracle_clip
    cfg = ClipTextConfig(num_hidden_layers=3)                    # 3 of the 12 identical layers keep the CPU test short
    o = make_oracle_clip(seed=2, cfg=cfg, round_to=None)
    tc = transformers.CLIPTextConfig(hidden_size=768, intermediate_size=3072, num_hidden_layers=3,
                                     num_attention_heads=12, max_position_embeddings=77, vocab_size=49408,
                                     hidden_act="quick_gelu")
    m = transformers.CLIPTextModel(tc).eval()
    res = m.load_state_dict(o.state_dict(), strict=False)
    assert not res.unexpected_keys and all(k.endswith("position_ids") for k in res.missing_keys)
    ids = torch.randint(0, 49408, (2, 77), generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        ref = m(ids)[0]
    out = o(ids)
    assert out.shape == (2, 77, 768)
    assert ((out - ref).norm() / ref.norm()).item() < 1e-5
    # causal: the embedding of token i does not depend on later tokens
    ids2 = ids.clone()
    ids2[:, 40:] = 7
    assert torch.allclose(o(ids2)[:, :40], out[:, :40], atol=1e-5)


def test_oracle_metrics_known_answers():
    """Row f4's oracle (the reference's formulas, 06:209-284): closed-form cases."""
    from oracle import metrics as OM
    a = torch.rand(3, 16, 20)
    assert OM.compute_mse(a, a) == 0.0 and OM.compute_psnr(0.0) == 100.0
    assert abs(OM.compute_psnr(0.01) - 20.0) < 1e-9
    ramp = torch.stack([torch.full((3, 8, 8), 0.1 * t) for t in range(5)])          # linear in time: no flicker
    assert OM.compute_flicker_index(ramp) < 1e-7
    osc = torch.stack([torch.full((3, 8, 8), float(t % 2)) for t in range(5)])      # 0,1,0,1,0: |0-2+0| = 2
    assert abs(OM.compute_flicker_index(osc) - 2.0) < 1e-7
    import numpy as np
    assert torch.allclose(OM.warp_frame(a, np.zeros((16, 20, 2), np.float32)), a, atol=1e-6)
    shift = np.zeros((16, 20, 2), np.float32)
    shift[..., 0] = 1.0                                                             # sample one pixel to the right
    w = OM.warp_frame(a, shift)
    assert torch.allclose(w[:, :, :-1], a[:, :, 1:], atol=1e-5) and torch.allclose(w[:, :, -1], a[:, :, -1], atol=1e-5)


def test_oracle_lpips_alexnet_geometry_matches_torchvision():
    tv = pytest.importorskip("torchvision")
    from oracle.lpips_alex import ALEX_CONVS, make_oracle_lpips, lpips_consecutive
    feats = tv.models.alexnet(weights=None).features
    for sl, idx, cin, cout, k, s, p, pool in ALEX_CONVS:
        c = feats[idx]
        assert (c.in_channels, c.out_channels, c.kernel_size[0], c.stride[0], c.padding[0]) == (cin, cout, k, s, p)
        assert isinstance(feats[idx + 1], torch.nn.ReLU) and (not pool or isinstance(feats[idx - 1], torch.nn.MaxPool2d))
    m = make_oracle_lpips()
    fr = torch.rand(3, 3, 64, 64, generator=torch.Generator().manual_seed(0))
    d = lpips_consecutive(m, fr)
    assert len(d) == 2 and all(v > 0 for v in d)
    assert lpips_consecutive(m, torch.stack([fr[0], fr[0]]))[0] == 0.0

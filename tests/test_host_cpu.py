"""CPU tests (-m "not gpu") of the host side: scheduler tables of the product against the oracle and the golden
fixtures, the C-ABI library (loads, exports every symbol of include/b200ad.h, fails loudly without a GPU), and the
diffusers-compatible import surface the reference uses (05_grid_search_ablation.py:27-28)."""
import ctypes
import json
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

import b200ad  # noqa: E402,F401
from b200ad import lib  # noqa: E402
from b200ad.schedulers import SD15_SCHEDULER_CONFIG, DDIMScheduler, EulerDiscreteScheduler  # noqa: E402


def _ddim():
    return DDIMScheduler.from_config(SD15_SCHEDULER_CONFIG, beta_schedule="linear", steps_offset=1, clip_sample=False)


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(GOLDEN, "schedulers.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("n", [15, 20, 25, 30, 40, 50])
def test_ddim_indexing_bit_exact_and_coefficients(golden, n):
    s = _ddim()
    s.set_timesteps(n)
    g = golden["ddim"][str(n)]
    assert s.timesteps.dtype == torch.int64
    assert s.timesteps.tolist() == g["timesteps"]                       # integer indexing: bit-exact
    assert s._prev_t[-1] == g["timesteps"][-1] - 1000 // n < 0          # last step falls back to final alpha
    for i in range(n):
        cx, ce, nxt = s.step_coefficients(i)
        assert abs(cx - g["cx"][i]) <= 1e-6 * abs(g["cx"][i])           # alphas_cumprod is an fp32 table
        assert abs(ce - g["ce"][i]) <= 1e-6 * max(abs(g["ce"][i]), 1e-3)
        assert nxt == 1.0


def test_ddim_config_round_trip_and_overrides():
    s = _ddim()
    c = s.config
    assert c.beta_schedule == "linear" and c.steps_offset == 1 and c.clip_sample is False
    assert c.set_alpha_to_one is False and c.beta_start == 0.00085 and c.beta_end == 0.012
    assert abs(float(s.alphas_cumprod[999]) - 1.578962780e-03) < 1e-9
    assert float(s.final_alpha_cumprod) == float(s.alphas_cumprod[0])
    s2 = DDIMScheduler.from_config(s.config)
    s2.set_timesteps(25)
    assert s2.timesteps.tolist()[:3] == [961, 921, 881]
    assert s.init_noise_sigma == 1.0 and s.order == 1
    x = torch.randn(4)
    assert s.scale_model_input(x, 961) is x
    with pytest.raises(NotImplementedError):
        DDIMScheduler.from_config(SD15_SCHEDULER_CONFIG, clip_sample=True)


@pytest.mark.parametrize("n", [15, 25, 50])
def test_euler_tables(golden, n):
    s = EulerDiscreteScheduler.from_config(SD15_SCHEDULER_CONFIG, timestep_spacing="linspace", beta_schedule="linear")
    s.set_timesteps(n)
    g = golden["euler"][str(n)]
    assert np.allclose(s.timesteps.numpy(), np.array(g["timesteps"], dtype=np.float32), rtol=0, atol=0)
    assert np.allclose(s.sigmas.numpy(), np.array(g["sigmas"], dtype=np.float32), rtol=1e-7, atol=0)
    assert abs(s.init_noise_sigma - max(g["sigmas"])) < 1e-6
    sg = np.array(g["sigmas"], dtype=np.float64)
    for i in range(n):
        cx, ce, nxt = s.step_coefficients(i)
        assert cx == 1.0 and abs(ce - (sg[i + 1] - sg[i])) < 1e-6
        assert abs(nxt - 1.0 / np.sqrt(sg[i + 1] ** 2 + 1)) < 1e-6
    assert abs(s.first_in_scale() - 1.0 / np.sqrt(sg[0] ** 2 + 1)) < 1e-7


def test_step_rejects_unknown_timestep_and_missing_set_timesteps():
    s = _ddim()
    with pytest.raises(ValueError):
        s.step(torch.zeros(1), 961, torch.zeros(1))
    s.set_timesteps(25)
    with pytest.raises(ValueError):
        s._step_index_of(960)


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "b200ad.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200ad_\w+)\s*\(", src)))


def test_cabi_library_exports_every_declared_symbol():
    syms = _header_symbols()
    assert len(syms) >= 18
    assert os.path.exists(lib.LIB_PATH), "libb200ad.so missing: run `python -c 'import __graft_entry__ as g; g.build()'`"
    L = ctypes.CDLL(lib.LIB_PATH)
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/b200ad.h but not exported"
    assert sorted(lib.SIGNATURES) == syms, "lib.py SIGNATURES and include/b200ad.h disagree"
    assert lib.load().b200ad_version() >= 100


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_anywhere():
    from b200ad.ops import get_ops
    from b200ad.pipeline import AnimateDiffPipeline
    from b200ad.unet import UNetMotionConfig, UNetMotionModel
    o = get_ops("fp32")
    with pytest.raises(lib.B200ADError):
        o.cfg_step(torch.zeros(4), torch.zeros(4), torch.zeros(4), 7.5, 1.0, 0.0)
    with pytest.raises(RuntimeError):
        AnimateDiffPipeline.from_pretrained("runwayml/stable-diffusion-v1-5")
    m = UNetMotionModel(UNetMotionConfig(block_out_channels=(32, 64, 64, 64), norm_num_groups=8,
                                         motion_norm_num_groups=8, num_attention_heads=2,
                                         motion_num_attention_heads=2, cross_attention_dim=48))
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 4, 4, 8, 8), 1, encoder_hidden_states=torch.zeros(1, 77, 48))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "video-diffusion-experiments_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f"{f} imports oracle/"


def test_product_module_tree_matches_oracle_state_dict_keys():
    from b200ad.unet import UNetMotionConfig, UNetMotionModel
    from oracle.unet_motion import UNetMotionConfig as OC, UNetMotionModel as OM
    with torch.device("meta"):
        a, b = UNetMotionModel(UNetMotionConfig()), OM(OC())
    ka, kb = a.state_dict(), b.state_dict()
    assert set(ka) == set(kb)
    assert all(ka[k].shape == kb[k].shape for k in ka)
    assert sum(p.numel() for p in a.parameters()) == 1_312_730_244
    names = dict(a.named_modules())
    attn = names["down_blocks.0.attentions.0.transformer_blocks.0.attn1"]
    assert type(attn).__name__ == "Attention" and attn.heads == 8 and attn.to_q.in_features == 320
    assert len(a.attn_processors) == 32 + 21 * 2


def test_diffusers_shim_import_surface():
    sys.path.insert(0, os.path.join(ROOT, "compat"))
    try:
        import diffusers
        from diffusers import AnimateDiffPipeline, DDIMScheduler as D, EulerDiscreteScheduler as E, MotionAdapter
        from diffusers.utils import export_to_gif
        assert callable(export_to_gif) and D is DDIMScheduler and E is EulerDiscreteScheduler
        assert hasattr(AnimateDiffPipeline, "from_pretrained") and hasattr(MotionAdapter, "from_pretrained")
        assert "b200ad" in diffusers.__version__
    finally:
        sys.path.remove(os.path.join(ROOT, "compat"))
        for k in [k for k in sys.modules if k == "diffusers" or k.startswith("diffusers.")]:
            del sys.modules[k]


def test_euler_init_noise_sigma_follows_the_current_schedule():
    """with SD1.5's own scheduler config (timestep_spacing='leading', steps_offset=1) Euler's
    init_noise_sigma depends on THIS call's sigma table, so the pipeline sets the timesteps before it draws / scales the
    latents (diffusers' order).  Known answers: 14.6488 before set_timesteps, 11.7188 at 25 steps."""
    from b200ad.schedulers import EulerDiscreteScheduler, SD15_SCHEDULER_CONFIG
    from oracle.schedulers import OracleEuler
    s = EulerDiscreteScheduler.from_config(SD15_SCHEDULER_CONFIG)
    assert abs(s.init_noise_sigma - 14.6488) < 1e-3
    for n in (25, 15, 25):                                   # consecutive calls with different step counts
        s.set_timesteps(n)
        o = OracleEuler(beta_schedule="scaled_linear", timestep_spacing="leading", steps_offset=1)
        o.set_timesteps(n)
        assert abs(s.init_noise_sigma - o.init_noise_sigma) < 1e-6
        assert abs(s.first_in_scale() - 1.0 / (float(o.sigmas[0]) ** 2 + 1) ** 0.5) < 1e-7
    s.set_timesteps(25)
    assert abs(s.init_noise_sigma - 11.7188) < 1e-3
    for sp in ("linspace", "trailing"):
        e = EulerDiscreteScheduler.from_config(SD15_SCHEDULER_CONFIG, timestep_spacing=sp)
        e.set_timesteps(25)
        assert abs(e.init_noise_sigma - float(e.sigmas.max())) < 1e-7
    import inspect
    from b200ad import pipeline
    src = inspect.getsource(pipeline.AnimateDiffPipeline.__call__)
    assert src.index("set_timesteps") < src.index("prepare_latents")


def test_checkpoint_loading_refuses_silent_partial_loads(tmp_path):
    """a key mismatch must raise instead of leaving seeded-random weights in place; SD1.5's legacy VAE
    attention names (query/key/value/proj_attn) are renamed on load."""
    import torch.nn as nn
    from b200ad.pipeline import _load_checked, _rename_legacy_vae_keys
    m = nn.Sequential(nn.Linear(4, 4), nn.Linear(4, 4))
    sd = m.state_dict()
    _load_checked(m, sd, "ok")
    bad = {k: v for k, v in sd.items() if not k.startswith("1.")}
    with pytest.raises(RuntimeError, match="missing"):
        _load_checked(m, bad, "partial")
    _load_checked(m, bad, "allowed", allow_missing=lambda k: k.startswith("1."))
    with pytest.raises(RuntimeError, match="unexpected"):
        _load_checked(m, {**sd, "2.weight": torch.zeros(1)}, "extra")
    legacy = {"decoder.mid_block.attentions.0.query.weight": torch.zeros(8, 8, 1, 1),
              "decoder.mid_block.attentions.0.proj_attn.bias": torch.zeros(8),
              "decoder.mid_block.attentions.0.group_norm.weight": torch.zeros(8)}
    out = _rename_legacy_vae_keys(legacy)
    assert set(out) == {"decoder.mid_block.attentions.0.to_q.weight", "decoder.mid_block.attentions.0.to_out.0.bias",
                        "decoder.mid_block.attentions.0.group_norm.weight"}
    assert out["decoder.mid_block.attentions.0.to_q.weight"].shape == (8, 8)


def test_denoise_plan_table_is_keyed_on_scheduler_config_contents():
    import inspect
    from b200ad import pipeline
    src = inspect.getsource(pipeline.DenoisePlan.load)
    assert "id(scheduler)" not in src and "scheduler.config" in src


def test_vae_state_dict_follows_diffusers_names():
    from b200ad.vae import AutoencoderKLDecoder
    from oracle.vae import make_oracle_vae
    with torch.device("meta"):
        v = AutoencoderKLDecoder()
    keys = set(v.state_dict())
    for k in ("post_quant_conv.weight", "decoder.conv_in.weight", "decoder.mid_block.attentions.0.to_q.bias",
              "decoder.mid_block.attentions.0.to_out.0.weight", "decoder.mid_block.resnets.1.conv2.weight",
              "decoder.up_blocks.0.upsamplers.0.conv.weight", "decoder.up_blocks.2.resnets.0.conv_shortcut.weight",
              "decoder.up_blocks.3.resnets.2.norm2.bias", "decoder.conv_norm_out.weight", "decoder.conv_out.bias"):
        assert k in keys, k
    assert keys == set(make_oracle_vae(up_channels=(512, 512, 256, 128)).state_dict())
    n = sum(p.numel() for p in v.parameters())
    assert n == 49_490_199, n            # SD1.5 AutoencoderKL decoder + post_quant_conv


def test_sweep_exporter_writes_the_reference_layout(tmp_path):
    """FrameExporter (row f3) in its synchronous form: frames/frame_%04d.png, <id>.gif, config.json last (05:172-190)."""
    import numpy as np
    from b200ad import sweep
    run = sweep.sweep_runs(video_filter="portrait", phase="cfg")[0]
    arr = (np.random.default_rng(0).random((3, 16, 16, 3)) * 255).astype(np.uint8)
    exp = tmp_path / run.experiment_id
    (exp / "frames").mkdir(parents=True)
    sweep._encode_run(arr, run, exp)
    assert sorted(p.name for p in (exp / "frames").iterdir()) == [f"frame_{i:04d}.png" for i in range(3)]
    assert (exp / f"{run.experiment_id}.gif").exists() and sweep.is_done(run, tmp_path)
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(exp / "frames" / "frame_0001.png")), arr[1])


def test_subpixel_upsample_weights_reproduce_upsample_then_conv():
    """unet.upconv_weights (what b200ad_upconv3x3 consumes): nearest x2 + conv3x3(pad 1) == four 2x2 convolutions of the
    low-resolution image, phase (py, px) reading source pixels (y + ty - 1 + py, x + tx - 1 + px) -- checked in fp64 on
    the CPU, borders included (the kernel's TMA zero fill = the zero padding of the upsampled image)."""
    import torch.nn.functional as F
    from b200ad.unet import upconv_weights
    g = torch.Generator().manual_seed(0)
    n, cin, cout, h, w = 2, 5, 3, 4, 6
    x = torch.randn(n, cin, h, w, generator=g, dtype=torch.float64)
    wt = torch.randn(cout, cin, 3, 3, generator=g, dtype=torch.float64)
    ref = F.conv2d(F.interpolate(x, scale_factor=2.0, mode="nearest"), wt, padding=1)          # [n, cout, 2h, 2w]
    w4 = upconv_weights(wt).reshape(4, cout, 2, 2, cin)
    xp = F.pad(x, (1, 1, 1, 1))                                                                # zero halo
    out = torch.zeros_like(ref)
    for py in (0, 1):
        for px in (0, 1):
            acc = torch.zeros(n, cout, h, w, dtype=torch.float64)
            for ty in (0, 1):
                for tx in (0, 1):
                    dy, dx = ty - 1 + py, tx - 1 + px
                    src = xp[:, :, 1 + dy:1 + dy + h, 1 + dx:1 + dx + w]
                    acc += torch.einsum("oc,nchw->nohw", w4[py * 2 + px, :, ty, tx], src)
            out[:, :, py::2, px::2] = acc
    assert (out - ref).abs().max() < 1e-12


def test_fold_tail_weights_equal_ff2_then_proj_out():
    """unet.fold_tail_weights: x + Wout (h + W2 g + b2) + bout == x + [h | g] [Wout | Wout W2]^T + (Wout b2 + bout)."""
    from b200ad.unet import fold_tail_weights
    g = torch.Generator().manual_seed(1)
    C, M = 8, 11
    r = lambda *s: torch.randn(*s, generator=g, dtype=torch.float64)
    w2, b2, wo, bo = r(C, 4 * C), r(C), r(C, C, 1, 1), r(C)                  # proj_out as a 1x1 conv weight
    x, h, gate = r(M, C), r(M, C), r(M, 4 * C)
    ref = x + (h + gate @ w2.T + b2) @ wo.reshape(C, C).T + bo
    wcat, bcat = fold_tail_weights(w2, b2, wo, bo)
    assert wcat.shape == (C, 5 * C)
    out = x + torch.cat([h, gate], 1) @ wcat.T + bcat
    assert (out - ref).abs().max() < 1e-12


def test_attention_head_stride(monkeypatch):
    """unet._dpad: the narrowest stride the attention kernels accept (a multiple of 8, >= 16*ceil(d/16), room for the
    ones column at index d); whole slabs with B200AD_TIGHT_HEADS=0; none in the fp32 mode."""
    from types import SimpleNamespace
    from b200ad.unet import _dpad
    bf, f32 = SimpleNamespace(mode="bf16"), SimpleNamespace(mode="fp32")
    monkeypatch.delenv("B200AD_TIGHT_HEADS", raising=False)
    assert [_dpad(d, bf) for d in (40, 80, 160, 64, 8)] == [48, 88, 168, 64, 16]
    assert _dpad(40, f32) == 40
    for d in (8, 24, 40, 64, 80, 160):
        p = _dpad(d, bf)
        assert p % 8 == 0 and p >= 16 * ((d + 15) // 16) and (p > d or p in (64, 128, 192))
    monkeypatch.setenv("B200AD_TIGHT_HEADS", "0")
    assert [_dpad(d, bf) for d in (40, 80, 160)] == [64, 128, 192]
    with pytest.raises(ValueError):
        _dpad(200, bf)

"""Hook-visible module mode (SURVEY.md 8b "introspection" row; -m gpu): with forward hooks registered -- what
/root/reference/utils/forward_tracer.py:177-194 does to every named module, driven by
/root/reference/experiments/03_trace_forward_pass.py:105-113 -- UNetMotionModel.forward runs through its sub-modules
with diffusers' tensor shapes; every module fires, the result matches the oracle and the fused plan."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SMALL = dict(block_out_channels=(64, 128, 256, 256))


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


def _pair():
    from oracle.unet_motion import UNetMotionConfig as OC, make_oracle_unet
    from b200ad.unet import UNetMotionConfig as PC, UNetMotionModel
    oracle = make_oracle_unet(seed=0, cfg=OC(**SMALL))
    prod = UNetMotionModel(PC(**SMALL))
    prod.load_state_dict(oracle.state_dict())
    return oracle, prod.cuda()


@pytest.mark.parametrize("mode,tol", [("bf16", 1e-2), ("fp32", 1e-4)])
def test_hooks_fire_on_every_module_and_result_matches(mode, tol):
    oracle, prod = _pair()
    prod.set_mode(mode)
    B, F, h, w = 2, 8, 16, 16
    lat = torch.randn(B, 4, F, h, w, generator=torch.Generator().manual_seed(42))
    txt = torch.randn(B, 77, 768, generator=torch.Generator().manual_seed(43))
    fused = prod(lat.cuda(), 500, encoder_hidden_states=txt.cuda()).sample
    records = {}

    def mk(name):
        def hook(mod, inp, out):
            t = inp[0] if inp and torch.is_tensor(inp[0]) else None
            records[name] = (type(mod).__name__, None if t is None else tuple(t.shape),
                             tuple(out.shape) if torch.is_tensor(out) else None)
        return hook
    handles = [m.register_forward_hook(mk(n)) for n, m in prod.named_modules() if n]
    out = prod(lat.cuda(), 500, encoder_hidden_states=txt.cuda()).sample
    for hd in handles:
        hd.remove()
    expected = {n for n, m in prod.named_modules() if n and not isinstance(m, torch.nn.ModuleList)}
    assert set(records) == expected, sorted(expected - set(records))[:8]
    ref = oracle(lat, 500, encoder_hidden_states=txt).sample
    e_ref, e_fused = rel(out.cpu(), ref), rel(out, fused)
    print(f"module mode {mode}: vs oracle {e_ref:.3e}, vs fused plan {e_fused:.3e}")
    assert e_ref < tol and e_fused < tol
    # diffusers' shapes at the boundaries the reference's tracer reports (03_trace_forward_pass.py:160-169,269-274)
    name = "down_blocks.0.motion_modules.0.transformer_blocks.0.attn1"
    assert records[name][0] == "Attention" and records[name][1] == (B * h * w, F, 64)
    assert records["down_blocks.0.motion_modules.0"][:2] == ("AnimateDiffTransformer3D", (B * F, 64, h, w))
    assert records["down_blocks.0.attentions.0.transformer_blocks.0.attn1"][1] == (B * F, h * w, 64)
    assert records["down_blocks.0.attentions.0.transformer_blocks.0.attn2.to_k"][:2] == ("Linear", (B * F, 77, 768))
    assert records["down_blocks.0.resnets.0.conv1"][0] == "Conv2d"
    assert records["mid_block.resnets.1.norm2"][0] == "GroupNorm"
    # hooks removed: the fused plan runs again (bitwise the same as before)
    assert torch.equal(prod(lat.cuda(), 500, encoder_hidden_states=txt.cuda()).sample, fused)


def test_module_mode_text_batch_conventions_and_fp16_io():
    """03_trace_forward_pass.py:86-113 calls the UNet with fp16 sample / timestep tensor / [B,77,768] text."""
    _, prod = _pair()
    prod.module_mode = True
    lat = torch.randn(1, 4, 4, 8, 8, generator=torch.Generator().manual_seed(1))
    txt = torch.randn(1, 77, 768, generator=torch.Generator().manual_seed(2))
    a = prod(lat.cuda().half(), torch.tensor([500], device="cuda", dtype=torch.float16),
             encoder_hidden_states=txt.cuda().half()).sample
    assert a.dtype == torch.float16 and a.shape == lat.shape
    prod.module_mode = False
    b = prod(lat.cuda().half(), 500, encoder_hidden_states=txt.cuda().half()).sample
    assert rel(a, b) < 3e-2            # fp16 activations between modules vs the fused plan's fp32 stream

"""VAE decode + postprocess (-m gpu): b200ad AutoencoderKLDecoder (sm_100a kernels through
the C ABI) against the CPU oracle (oracle/vae.py) on identical bf16-representable random-init weights.
The frames are what /root/reference/experiments/05_grid_search_ablation.py:169,178-179 consumes."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SMALL = (128, 128, 64, 64)


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


def _pair(chans, mode="bf16"):
    from oracle.vae import make_oracle_vae
    from b200ad.vae import AutoencoderKLDecoder
    oracle = make_oracle_vae(seed=1, up_channels=chans)
    prod = AutoencoderKLDecoder(chans, mode=mode)
    prod.load_state_dict(oracle.state_dict())
    return oracle, prod.cuda()


@pytest.mark.parametrize("mode,tol", [("bf16", 2e-2), ("fp32", 1e-4)])
@pytest.mark.parametrize("F,h,w", [(4, 8, 8), (3, 16, 8), (2, 32, 32)])
def test_small_vae_matches_oracle(mode, tol, F, h, w):
    from oracle.vae import decode_latents, postprocess_video, SCALING_FACTOR
    oracle, prod = _pair(SMALL, mode)
    lat = torch.randn(1, 4, F, h, w, generator=torch.Generator().manual_seed(3)) * 0.8
    ref = decode_latents(oracle, lat)
    out = prod.decode_video(lat.cuda())
    assert out.shape == ref.shape == (1, 3, F, 8 * h, 8 * w)
    e = rel(out.cpu(), ref)
    print(f"small vae {mode} F={F} {h}x{w}: rel-L2 {e:.3e}")
    assert e < tol
    # diffusers' AutoencoderKL.decode convention: z already divided by the scaling factor, [N,4,h,w] -> [N,3,H,W]
    z = (lat / SCALING_FACTOR)[0].permute(1, 0, 2, 3).contiguous()
    img = prod.decode(z.cuda())
    assert rel(img.cpu(), ref[0].permute(1, 0, 2, 3)) < tol
    # uint8 frames: at most one grey level off (rounding of an already-rounded value) in fp32 mode
    u8 = prod.decode_video_u8(lat.cuda()).cpu()
    ref8 = postprocess_video(ref)
    assert u8.shape == ref8.shape == (F, 8 * h, 8 * w, 3) and u8.dtype == torch.uint8
    diff = (u8.int() - ref8.int()).abs()
    if mode == "fp32":
        assert diff.max() <= 1 and (diff > 0).float().mean() < 1e-2
    else:
        assert diff.float().mean() < 1.5


def test_vae_slicing_and_module_mode_hooks():
    """enable_vae_slicing (05_grid_search_ablation.py:144) must not change the result; with forward hooks registered the
    decoder runs through its sub-modules (same numbers within bf16 rounding) and every hooked module fires."""
    from oracle.vae import decode_latents, SCALING_FACTOR
    oracle, prod = _pair(SMALL, "bf16")
    lat = torch.randn(1, 4, 3, 8, 8, generator=torch.Generator().manual_seed(4)) * 0.8
    base = prod.decode_video(lat.cuda())
    prod.use_slicing = True
    assert torch.equal(prod.decode_video(lat.cuda()), base)
    prod.use_slicing = False
    fired = set()
    handles = [m.register_forward_hook(lambda mod, i, o, n=n: fired.add(n)) for n, m in prod.named_modules()
               if n and not isinstance(m, torch.nn.ModuleList)]
    z = (lat / SCALING_FACTOR)[0].permute(1, 0, 2, 3).contiguous()
    img = prod.decode(z.cuda())
    for hd in handles:
        hd.remove()
    expected = {n for n, m in prod.named_modules() if n and not isinstance(m, torch.nn.ModuleList)}
    assert fired == expected, sorted(expected - fired)[:5]
    ref = decode_latents(oracle, lat)[0].permute(1, 0, 2, 3)
    assert rel(img.cpu(), ref) < 2e-2


def test_full_vae_256_matches_oracle():
    """SD1.5 decoder geometry (512, 512, 256, 128), 4 frames of 256x256 (latents 32x32): the conv / GroupNorm / attention
    kernels at the VAE's own shapes (C = 128 at 256^2 pixels, one attention head of d = 512 over 1024 tokens)."""
    from oracle.vae import decode_latents, postprocess_video, psnr
    oracle, prod = _pair((512, 512, 256, 128), "bf16")
    lat = torch.randn(1, 4, 4, 32, 32, generator=torch.Generator().manual_seed(5)) * 0.8
    ref = decode_latents(oracle, lat)
    out = prod.decode_video(lat.cuda())
    e = rel(out.cpu(), ref)
    p = psnr(postprocess_video(out.cpu()).float() / 255, postprocess_video(ref).float() / 255)
    print(f"full vae bf16 4f 256^2: rel-L2 {e:.3e}, PSNR of the uint8 frames {p:.1f} dB")
    assert e < 2e-2 and p > 35.0
    prod.set_mode("fp32")
    e32 = rel(prod.decode_video(lat.cuda()).cpu(), ref)
    print(f"full vae fp32 mode: rel-L2 {e32:.3e}")
    assert e32 < 1e-4

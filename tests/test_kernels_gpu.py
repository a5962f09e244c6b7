"""Kernel-level parity (-m gpu): each C-ABI entry point against a plain PyTorch fp32 reference of the same op
on identical seeded inputs.  bf16 mode: inputs are made bf16-representable so the only differences are
accumulation order and the single output rounding; tolerances are written per test.  fp32 mode: <= 1e-5."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda"


def ops(mode):
    import b200ad.ops as O
    return O.get_ops(mode)


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale)


def bf(x):
    return x.to(torch.bfloat16).float()


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-20)).item()


MODES = [("bf16", 6e-3), ("fp32", 2e-5)]


@pytest.mark.parametrize("mode,tol", MODES)
@pytest.mark.parametrize("M,N,K", [(256, 128, 64), (1000, 320, 320), (4096, 1280, 640), (300, 2560, 320),
                                   (128, 32, 2880), (77, 640, 768)])
def test_gemm_bias_residual(mode, tol, M, N, K):
    o = ops(mode)
    a, w = bf(rnd(M, K, seed=1)), bf(rnd(N, K, seed=2, scale=K ** -0.5))
    bias, res = rnd(N, seed=3), rnd(M, N, seed=4)
    ref = a @ w.t() + bias + res
    out32, outop = o.gemm(a.to(DEV, o.op_dtype), w.to(DEV, o.op_dtype), bias=bias.to(DEV), residual=res.to(DEV),
                          want_f32=True, want_op=True)
    assert rel(out32.cpu(), ref) < (1e-5 if mode == "fp32" else 1e-5 + 2e-6)   # fp32 accumulate of exact bf16 products
    assert rel(outop.cpu(), ref) < tol


@pytest.mark.parametrize("mode,tol", MODES)
def test_gemm_inplace_residual_and_rowbias(mode, tol):
    o = ops(mode)
    M, N, K = 512, 320, 320
    a, w = bf(rnd(M, K, seed=1)), bf(rnd(N, K, seed=2, scale=K ** -0.5))
    rb = rnd(4, N, seed=5)
    x = rnd(M, N, seed=6)
    ref = a @ w.t() + rb.repeat_interleave(M // 4, 0) + x
    xd = x.to(DEV)
    o.gemm(a.to(DEV, o.op_dtype), w.to(DEV, o.op_dtype), row_bias=rb.to(DEV), rb_div=M // 4, residual=xd, out_f32=xd)
    assert rel(xd.cpu(), ref) < 2e-5


@pytest.mark.parametrize("mode,tol", MODES)
def test_gemm_geglu(mode, tol):
    o = ops(mode)
    M, C = 384, 320
    a = bf(rnd(M, C, seed=1))
    w = bf(rnd(8 * C, C, seed=2, scale=C ** -0.5))
    b = rnd(8 * C, seed=3)
    hg = a @ w.t() + b
    h, g = hg.chunk(2, -1)
    ref = h * F.gelu(g)
    wi = torch.stack([w[:4 * C], w[4 * C:]], 1).reshape(8 * C, C).contiguous()
    bi = torch.stack([b[:4 * C], b[4 * C:]], 1).reshape(8 * C).contiguous()
    _, out = o.gemm(a.to(DEV, o.op_dtype), wi.to(DEV, o.op_dtype), bias=bi.to(DEV), want_op=True, geglu=True)
    assert out.shape == (M, 4 * C)
    assert rel(out.cpu(), ref) < tol


def _conv_ref(x_nhwc, w, b, stride=1):
    y = F.conv2d(x_nhwc.permute(0, 3, 1, 2), w, b, stride=stride, padding=1)
    return y.permute(0, 2, 3, 1).contiguous()


def _pack_conv(w):
    return w.permute(0, 2, 3, 1).reshape(w.shape[0], -1).contiguous()


@pytest.mark.parametrize("mode,tol", MODES)
@pytest.mark.parametrize("n,h,w,cin,cout", [(2, 64, 64, 64, 64), (3, 32, 32, 128, 320), (5, 16, 16, 320, 64),
                                            (6, 8, 8, 64, 128), (9, 4, 4, 64, 64), (2, 12, 12, 64, 64),
                                            (1, 24, 40, 64, 96)])
def test_conv3x3(mode, tol, n, h, w, cin, cout):
    o = ops(mode)
    x = bf(rnd(n, h, w, cin, seed=1))
    wt = bf(rnd(cout, cin, 3, 3, seed=2, scale=(9 * cin) ** -0.5))
    b = rnd(cout, seed=3)
    temb = rnd(n, cout, seed=4)
    ref = _conv_ref(x, wt, b) + temb[:, None, None, :]
    out, _ = o.conv3x3(x.to(DEV, o.op_dtype), _pack_conv(wt).to(DEV, o.op_dtype), n, h, w, bias=b.to(DEV),
                       row_bias=temb.to(DEV), rb_div=h * w)
    assert rel(out.cpu().view(n, h, w, cout), ref) < 2e-5


@pytest.mark.parametrize("mode,tol", MODES)
@pytest.mark.parametrize("n,h,w,c", [(2, 32, 32, 64), (3, 8, 8, 128), (1, 64, 64, 64)])
def test_conv3x3_stride2_and_upsample(mode, tol, n, h, w, c):
    o = ops(mode)
    x = bf(rnd(n, h, w, c, seed=1))
    wt = bf(rnd(c, c, 3, 3, seed=2, scale=(9 * c) ** -0.5))
    b = rnd(c, seed=3)
    xd = x.to(DEV)
    # Downsample2D: conv stride 2 pad 1
    ref = _conv_ref(x, wt, b, stride=2)
    s2d = o.cast(xd, n, h, w, c, mode=1)
    out, _ = o.conv3x3(s2d, _pack_conv(wt).to(DEV, o.op_dtype), n, h // 2, w // 2, stride2=True, bias=b.to(DEV))
    assert rel(out.cpu().view(n, h // 2, w // 2, c), ref) < 2e-5
    # Upsample2D: nearest x2 then conv
    up = F.interpolate(x.permute(0, 3, 1, 2), scale_factor=2.0, mode="nearest").permute(0, 2, 3, 1).contiguous()
    ref = _conv_ref(up, wt, b)
    upd = o.cast(xd, n, h, w, c, mode=2)
    out, _ = o.conv3x3(upd, _pack_conv(wt).to(DEV, o.op_dtype), n, 2 * h, 2 * w, bias=b.to(DEV))
    assert rel(out.cpu().view(n, 2 * h, 2 * w, c), ref) < 2e-5


@pytest.mark.parametrize("mode,tol", MODES)
@pytest.mark.parametrize("nimg,hw,c0,c1,fps,silu", [(4, 256, 320, 0, 1, True), (4, 64, 1280, 640, 1, True),
                                                    (8, 1024, 640, 320, 1, True), (8, 100, 320, 0, 4, False),
                                                    (2, 4096, 64, 0, 2, False), (3, 16, 2560, 0, 1, True),
                                                    # > 72 MB of fp32 input: walked in L2-sized sub-batches
                                                    (28, 1024, 640, 0, 1, True), (32, 1024, 320, 320, 16, False)])
def test_group_norm(mode, tol, nimg, hw, c0, c1, fps, silu):
    o = ops(mode)
    x0 = rnd(nimg * hw, c0, seed=1) * 2 + 0.5
    x1 = rnd(nimg * hw, c1, seed=2) if c1 else None
    C = c0 + c1
    gamma, beta = rnd(C, seed=3), rnd(C, seed=4)
    x = x0 if x1 is None else torch.cat([x0, x1], -1)
    xr = x.view(nimg // fps, fps * hw, C).permute(0, 2, 1)           # [stat, C, fps*hw]
    ref = F.group_norm(xr, 32, gamma, beta, eps=1e-6).permute(0, 2, 1).reshape(nimg * hw, C)
    if silu:
        ref = F.silu(ref)
    out, raw = o.group_norm(x0.to(DEV), None if x1 is None else x1.to(DEV), nimg, hw, 32, gamma.to(DEV),
                            beta.to(DEV), 1e-6, silu=silu, frames_per_stat=fps, want_raw=True)
    assert rel(out.cpu(), ref) < (4e-3 if mode == "bf16" else 1e-5)
    assert rel(raw.cpu(), x) < (4e-3 if mode == "bf16" else 1e-7)


@pytest.mark.parametrize("mode", ["bf16", "fp32"])
@pytest.mark.parametrize("c0,c1", [(640, 320), (320, 0), (1280, 640)])
def test_group_norm_is_deterministic_across_launches(mode, c0, c1):
    """The statistics kernel's last block of an image finalizes it (ticket counter).  With C = 960 the block has 240
    threads: the trailing half warp used to join that reduction and race warp 0 for four groups' (mean, rstd) -- one
    image in a few thousand came out wrong, which made the full-architecture parity tests flaky.  Many launches, all
    bitwise equal to the first and equal to the reference."""
    o = ops(mode)
    nimg, hw, C = 32, 256, c0 + c1
    x0 = (rnd(nimg * hw, c0, seed=1) * 2 + 0.5).to(DEV)
    x1 = rnd(nimg * hw, c1, seed=2).to(DEV) if c1 else None
    gamma, beta = rnd(C, seed=3).to(DEV), rnd(C, seed=4).to(DEV)
    x = x0 if x1 is None else torch.cat([x0, x1], -1)
    ref = F.silu(F.group_norm(x.cpu().view(nimg, hw, C).permute(0, 2, 1), 32, gamma.cpu(), beta.cpu(), eps=1e-5)
                 .permute(0, 2, 1).reshape(nimg * hw, C))
    first = o.group_norm(x0, x1, nimg, hw, 32, gamma, beta, 1e-5, silu=True)
    assert rel(first.cpu(), ref) < (4e-3 if mode == "bf16" else 1e-5)
    bad = 0
    for _ in range(300):
        out = o.group_norm(x0, x1, nimg, hw, 32, gamma, beta, 1e-5, silu=True)
        bad += int(not torch.equal(out, first))
    assert bad == 0, f"{bad} of 300 launches differ from the first"


@pytest.mark.parametrize("n,h,w,cin,sc_cin,cout", [(2, 32, 32, 64, 128, 64), (3, 8, 8, 128, 64, 320),
                                                   (5, 16, 16, 320, 960, 320), (2, 12, 20, 64, 64, 64),
                                                   (1, 64, 64, 64, 192, 128)])
def test_conv3x3_with_fused_1x1_shortcut(n, h, w, cin, sc_cin, cout):
    """b200ad_conv3x3_sc: conv3x3(a) + conv1x1(raw) + bias in one launch (the 1x1 is a second K segment of the implicit
    GEMM) against the two convolutions in fp32; plus its GroupNorm column statistics."""
    o = ops("bf16")
    a = bf(rnd(n, h, w, cin, seed=1))
    raw = bf(rnd(n, h, w, sc_cin, seed=2))
    w3 = bf(rnd(cout, cin, 3, 3, seed=3, scale=(9 * cin) ** -0.5))
    w1 = bf(rnd(cout, sc_cin, 1, 1, seed=4, scale=sc_cin ** -0.5))
    b = rnd(cout, seed=5) + 0.2
    ref = (_conv_ref(a, w3, b) + F.conv2d(raw.permute(0, 3, 1, 2), w1).permute(0, 2, 3, 1)).reshape(n * h * w, cout)
    wcat = torch.cat([_pack_conv(w3), w1.reshape(cout, sc_cin)], 1).contiguous().to(DEV, o.op_dtype)
    n0 = lib_launches()
    out = o.conv3x3_sc(a.to(DEV, o.op_dtype).view(n * h * w, cin), wcat, raw.to(DEV, o.op_dtype).view(n * h * w, sc_cin),
                       n, h, w, bias=b.to(DEV), stats=True)
    assert lib_launches() - n0 == 1
    assert rel(out.cpu(), ref) < 2e-5
    supported = o._stats_ok(True, h, w) > 0
    assert (out.data_ptr() in o.colstats) == supported
    if supported:
        st, spi = o.colstats[out.data_ptr()][:2]
        tot = st.view(n, spi, cout, 2).sum(1).cpu()
        per_img = ref.view(n, h * w, cout)
        assert (tot[..., 0] - per_img.sum(1)).abs().max() < 2e-3 * (h * w) ** 0.5
        assert rel(tot[..., 1], (per_img ** 2).sum(1)) < 1e-5


@pytest.mark.parametrize("n,h,w,cin,cout,exact", [(2, 32, 32, 64, 64, True), (3, 8, 8, 128, 320, True),
                                                  (5, 16, 16, 64, 128, False), (2, 12, 20, 64, 64, False),
                                                  (1, 64, 64, 64, 320, False), (4, 8, 8, 320, 64, True)])
def test_upconv3x3_subpixel_matches_upsample_then_conv(n, h, w, cin, cout, exact):
    """b200ad_upconv3x3 (nearest x2 + conv3x3 as four 2x2 sub-pixel convolutions over the low-resolution image) against
    F.interpolate(nearest) + conv2d in fp32.  exact=True: weights are multiples of 1/64 in [-1/2, 1/2], so the summed
    sub-pixel taps are bf16-exact and only the accumulation order differs (2e-5); otherwise the summed taps are rounded
    to bf16 once more than the direct form's weights (4e-3).  Also: the GroupNorm column statistics of the four phases
    cover the whole output image, and the GroupNorm that consumes them needs no statistics pass."""
    from b200ad.unet import _pack_upconv3x3
    o = ops("bf16")
    x = bf(rnd(n, h, w, cin, seed=1))
    if exact:
        g = torch.Generator().manual_seed(2)
        wt = torch.randint(-32, 33, (cout, cin, 3, 3), generator=g).float() / 64.0
    else:
        wt = bf(rnd(cout, cin, 3, 3, seed=2, scale=(9 * cin) ** -0.5))
    b = rnd(cout, seed=3) + 0.3
    up = F.interpolate(x.permute(0, 3, 1, 2), scale_factor=2.0, mode="nearest").permute(0, 2, 3, 1).contiguous()
    ref = _conv_ref(up, wt, b).reshape(n * 4 * h * w, cout)
    w4 = _pack_upconv3x3(o, wt)
    assert w4.shape == (4, cout, 4 * cin)
    n0 = lib_launches()
    out = o.upconv3x3(x.to(DEV, o.op_dtype), w4, n, h, w, bias=b.to(DEV), stats=True)
    assert lib_launches() - n0 == 1                                      # all four phases in one launch
    assert rel(out.cpu(), ref) < (2e-5 if exact else 4e-3)
    supported = o._stats_ok(True, h, w) > 0
    assert (out.data_ptr() in o.colstats) == supported
    gamma, beta = rnd(cout, seed=5), rnd(cout, seed=6)
    gref = F.silu(F.group_norm(out.cpu().view(n, 4 * h * w, cout).permute(0, 2, 1), 32, gamma, beta, eps=1e-5)
                  .permute(0, 2, 1).reshape(n * 4 * h * w, cout))
    n0 = lib_launches()
    gn = o.group_norm(out, None, n, 4 * h * w, 32, gamma.to(DEV), beta.to(DEV), 1e-5, silu=True)
    if supported:
        assert lib_launches() - n0 == 2                                  # finalize + apply
        st, spi = o.colstats[out.data_ptr()][:2]
        tot = st.view(n, spi, cout, 2).sum(1).cpu()
        per_img = out.cpu().view(n, 4 * h * w, cout)
        assert (tot[..., 0] - per_img.sum(1)).abs().max() < 2e-3 * (4 * h * w) ** 0.5
        assert rel(tot[..., 1], (per_img ** 2).sum(1)) < 1e-5
    assert rel(gn.cpu(), gref) < 4e-3


@pytest.mark.parametrize("nimg,h,w,cin,cout,fps", [(4, 16, 16, 64, 320, 1), (6, 8, 8, 64, 128, 1), (8, 8, 8, 128, 320, 4),
                                                   (3, 12, 12, 64, 64, 1), (2, 24, 40, 64, 128, 1), (4, 32, 32, 64, 640, 2),
                                                   (5, 8, 8, 64, 64, 1), (2, 64, 64, 64, 320, 1)])
def test_group_norm_from_conv_epilogue_statistics(nimg, h, w, cin, cout, fps):
    """GroupNorm whose statistics come from the producing conv's epilogue (b200ad_conv3x3_stats ->
    b200ad_gn_apply_stats): the GroupNorm reads its input once.  Geometries: several images per tile (8x8), boxes that
    overhang the image (12x12 rows, 24x40), an image count that does not fill the last tile, per-video statistics."""
    o = ops("bf16")
    x = bf(rnd(nimg, h, w, cin, seed=1))
    wt = bf(rnd(cout, cin, 3, 3, seed=2, scale=(9 * cin) ** -0.5))
    b = rnd(cout, seed=3) + 0.7
    res = rnd(nimg * h * w, cout, seed=4)
    conv = _conv_ref(x, wt, b).reshape(nimg * h * w, cout) + res
    gamma, beta = rnd(cout, seed=5), rnd(cout, seed=6)
    xr = conv.view(nimg // fps, fps * h * w, cout).permute(0, 2, 1)
    ref = F.silu(F.group_norm(xr, 32, gamma, beta, eps=1e-5).permute(0, 2, 1).reshape(nimg * h * w, cout))
    out32, _ = o.conv3x3(x.to(DEV, o.op_dtype), _pack_conv(wt).to(DEV, o.op_dtype), nimg, h, w, bias=b.to(DEV),
                         residual=res.to(DEV), stats=True)
    assert rel(out32.cpu(), conv) < 2e-5
    supported = o._stats_ok(True, h, w) > 0          # a tile must hold >= 32 pixels of one image (not 12x12: 4x4 boxes)
    assert (out32.data_ptr() in o.colstats) == supported, "column statistics emitted / not emitted unexpectedly"
    assert supported or (h, w) == (12, 12)
    n0 = lib_launches()
    out = o.group_norm(out32, None, nimg, h * w, 32, gamma.to(DEV), beta.to(DEV), 1e-5, silu=True, frames_per_stat=fps)
    if supported:
        assert lib_launches() - n0 == 2 + (fps > 1)                      # finalize (+ frame combine) + apply: no statistics pass
        st, spi = o.colstats[out32.data_ptr()][:2]
        tot = st.view(nimg, spi, cout, 2).sum(1).cpu()                   # per (image, channel) sum / sum of squares
        per_img = conv.view(nimg, h * w, cout)
        assert (tot[..., 0] - per_img.sum(1)).abs().max() < 2e-3 * (h * w) ** 0.5
        assert rel(tot[..., 1], (per_img ** 2).sum(1)) < 1e-5
    assert rel(out.cpu(), ref) < 4e-3
    # the same tensor through the two-pass path gives the same answer
    o.colstats.clear()
    out2 = o.group_norm(out32, None, nimg, h * w, 32, gamma.to(DEV), beta.to(DEV), 1e-5, silu=True, frames_per_stat=fps)
    assert rel(out.float(), out2.float()) < 2e-3


@pytest.mark.parametrize("nimg,hw,c0,c1,K", [(4, 256, 320, 0, 320), (2, 1024, 640, 320, 64), (8, 64, 1280, 640, 128),
                                             (16, 32, 320, 320, 64), (3, 4096, 320, 0, 64)])
def test_group_norm_from_gemm_epilogue_statistics_concat(nimg, hw, c0, c1, K):
    """Dense GEMM producers (proj_out of the transformers) and the UNet's skip concat: two tensors, each with its own
    epilogue statistics, normalised as one channel-concatenated input (group boundaries may straddle the two)."""
    o = ops("bf16")
    M = nimg * hw
    srcs = []
    for i, c in enumerate((c0, c1)):
        if not c:
            srcs.append((None, None))
            continue
        a, w = bf(rnd(M, K, seed=10 + i)), bf(rnd(c, K, seed=20 + i, scale=K ** -0.5))
        bias, res = rnd(c, seed=30 + i), rnd(M, c, seed=40 + i) * 2
        x = res.to(DEV)
        o.gemm(a.to(DEV, o.op_dtype), w.to(DEV, o.op_dtype), bias=bias.to(DEV), residual=x, out_f32=x, stats_hw=hw)
        assert x.data_ptr() in o.colstats
        srcs.append((x, a @ w.t() + bias + res))
    C = c0 + c1
    full = srcs[0][1] if not c1 else torch.cat([srcs[0][1], srcs[1][1]], -1)
    gamma, beta = rnd(C, seed=3), rnd(C, seed=4)
    ref = F.group_norm(full.view(nimg, hw, C).permute(0, 2, 1), 32, gamma, beta, eps=1e-6).permute(0, 2, 1).reshape(M, C)
    n0 = lib_launches()
    out, raw = o.group_norm(srcs[0][0], srcs[1][0], nimg, hw, 32, gamma.to(DEV), beta.to(DEV), 1e-6, silu=False,
                            want_raw=True)
    assert lib_launches() - n0 == 2
    assert rel(out.cpu(), ref) < 4e-3 and rel(raw.cpu(), full) < 4e-3
    # an in-place writer WITHOUT statistics invalidates the entry: the next GroupNorm takes the two-pass path
    x0 = srcs[0][0]
    o.gemm(bf(rnd(M, K, seed=50)).to(DEV, o.op_dtype), bf(rnd(c0, K, seed=51)).to(DEV, o.op_dtype), residual=x0, out_f32=x0)
    assert x0.data_ptr() not in o.colstats


def lib_launches():
    from b200ad import lib
    return lib.launch_count()


@pytest.mark.parametrize("mode,tol", MODES)
@pytest.mark.parametrize("rows,C,pe", [(1000, 320, False), (2 * 4 * 50, 640, True), (333, 1280, True), (64, 64, False)])
def test_layer_norm(mode, tol, rows, C, pe):
    o = ops(mode)
    x = rnd(rows, C, seed=1) * 3 + 1
    gamma, beta = rnd(C, seed=2), rnd(C, seed=3)
    ref = F.layer_norm(x, (C,), gamma, beta, 1e-5)
    hw, frames = 1, 1
    ped = None
    if pe:
        frames, hw = 4, 50 if rows == 400 else 1
        if rows == 333:
            frames, hw = 32, 3
        table = rnd(32, C, seed=4)
        f_idx = (torch.arange(rows) // hw) % frames
        ref = ref + table[f_idx]
        ped = table.to(DEV)
    out = o.layer_norm(x.to(DEV), gamma.to(DEV), beta.to(DEV), 1e-5, pe=ped, hw=hw, frames=frames)
    assert rel(out.cpu(), ref) < (4e-3 if mode == "bf16" else 1e-5)


def _attn_ref(q, k, v):
    return F.scaled_dot_product_attention(q, k, v)


@pytest.mark.parametrize("mode,tol", [("bf16", 8e-3), ("fp32", 2e-5)])
@pytest.mark.parametrize("nb,heads,Lq,Lk,d,dpad,kv_div", [
    (2, 8, 256, 256, 40, 64, 1), (3, 2, 1024, 1024, 40, 64, 1), (2, 8, 64, 64, 160, 192, 1),
    (2, 4, 256, 256, 80, 128, 1), (4, 8, 256, 77, 40, 64, 2), (2, 8, 64, 77, 160, 192, 2),
    (1, 2, 200, 300, 8, 64, 1), (2, 8, 1024, 77, 80, 128, 1), (2, 2, 1024, 1024, 80, 128, 1),
    (1, 2, 512, 512, 40, 64, 1), (1, 1, 2304, 2304, 80, 128, 1), (1, 2, 768, 768, 24, 64, 1),
    # benchmarked lengths without the V ones column (the in-kernel row-sum path): config 2 and config 5 level 0
    (2, 2, 4096, 4096, 40, 64, 1), (1, 1, 9216, 9216, 40, 64, 1),
    # single key tile (persistent cross-attention kernel): ragged query count, key counts 16 / 50 / 128, many tiles per CTA
    (3, 2, 200, 50, 40, 64, 1), (2, 4, 384, 128, 80, 128, 2), (2, 2, 640, 16, 160, 192, 1), (16, 8, 4096, 77, 40, 64, 16),
    # tight head strides (48 / 88 / 168 instead of whole 64-column slabs: a slab then runs into the next head, whose
    # columns no MMA may use) on every kernel: ping-pong, single-key-tile, short-sequence
    (2, 8, 256, 256, 40, 48, 1), (3, 2, 1024, 1024, 40, 48, 1), (2, 8, 64, 64, 160, 168, 1), (2, 4, 256, 256, 80, 88, 1),
    (4, 8, 256, 77, 40, 48, 2), (2, 8, 64, 77, 160, 168, 2), (2, 8, 1024, 77, 80, 88, 1), (2, 2, 1024, 1024, 80, 88, 1),
    (2, 3, 4096, 4096, 40, 48, 1), (1, 2, 200, 300, 8, 16, 1), (2, 8, 256, 256, 160, 160, 1)])
def test_attention(mode, tol, nb, heads, Lq, Lk, d, dpad, kv_div):
    o = ops(mode)
    nkv = nb // kv_div
    q = bf(rnd(nb, heads, Lq, d, seed=1))
    k = bf(rnd(nkv, heads, Lk, d, seed=2))
    v = bf(rnd(nkv, heads, Lk, d, seed=3))
    ref = _attn_ref(q, k.repeat_interleave(kv_div, 0), v.repeat_interleave(kv_div, 0))   # [nb, heads, Lq, d]
    ref = ref.permute(0, 2, 1, 3).reshape(nb * Lq, heads * d)

    def pack(t, L, n):   # [n, heads, L, d] -> [n*L, heads*dpad] zero padded
        p = torch.zeros(n, L, heads, dpad)
        p[..., :d] = t.permute(0, 2, 1, 3)
        return p.reshape(n * L, heads * dpad)

    qd = pack(q, Lq, nb).to(DEV, o.op_dtype)
    kvd = torch.cat([pack(k, Lk, nkv), pack(v, Lk, nkv)], -1).to(DEV, o.op_dtype)   # fused (k|v) buffer
    out = o.attention(qd, kvd[:, :heads * dpad], kvd[:, heads * dpad:], nb, heads, Lq, Lk, d, dpad, kv_div)
    assert rel(out.cpu(), ref) < tol


@pytest.mark.parametrize("nb,heads,L,d,dpad,grow", [(2, 3, 1024, 40, 64, False), (1, 2, 2304, 80, 128, False),
                                                    (1, 2, 1024, 40, 64, True), (2, 8, 256, 40, 64, False),
                                                    # the exact instance of the bench (config 2 level 0: L = 4096, d = 40,
                                                    # three query tiles per CTA, ragged last CTA) and config 5 (L = 9216)
                                                    (4, 8, 4096, 40, 64, False), (1, 2, 9216, 40, 64, False),
                                                    (1, 4, 4096, 80, 128, False),
                                                    # tight head strides: the stride the UNet uses (unet._dpad)
                                                    (4, 8, 4096, 40, 48, False), (1, 3, 2304, 80, 88, False),
                                                    (1, 2, 1024, 40, 48, True), (2, 8, 256, 40, 48, False)])
def test_attention_denominator_from_ones_column(nb, heads, L, d, dpad, grow):
    """B200AD_ATTN_V_ONES: V's pad column d holds 1.0, the long self-attention kernel takes the softmax denominator
    from the PV product (incl. across lazy rescales); the short-sequence kernel (L = 256) must simply ignore it."""
    o = ops("bf16")
    q = bf(rnd(nb, heads, L, d, seed=21) * (3.0 if grow else 1.0))
    k = bf(rnd(nb, heads, L, d, seed=22))
    if grow:
        k = bf(k * (1.0 + 3.0 * torch.arange(L)[None, None, :, None] / L))
    v = bf(rnd(nb, heads, L, d, seed=23))
    ref = _attn_ref(q, k, v).permute(0, 2, 1, 3).reshape(nb * L, heads * d)

    def pack(t, ones=False):
        p = torch.zeros(nb, L, heads, dpad)
        p[..., :d] = t.permute(0, 2, 1, 3)
        if ones:
            p[..., d] = 1.0
        return p.reshape(nb * L, heads * dpad).to(DEV, torch.bfloat16)

    out = o.attention(pack(q), pack(k), pack(v, ones=True), nb, heads, L, L, d, dpad, 1, v_ones=True)
    assert rel(out.cpu(), ref) < (1e-2 if grow else 8e-3)


@pytest.mark.parametrize("d,dpad", [(40, 64), (80, 128)])
def test_attention_lazy_rescale_with_growing_maxima(d, dpad):
    """Ping-pong kernel (attn_fa.cu): scores with a wide range and maxima that keep growing along the key axis, so
    the lazy O rescale (threshold 2^8) fires several times; compared with fp32 SDPA."""
    o = ops("bf16")
    nb, heads, L = 1, 2, 1024
    q = bf(rnd(nb, heads, L, d, seed=11) * 3.0)
    k = bf(rnd(nb, heads, L, d, seed=12) * (1.0 + 3.0 * torch.arange(L)[None, None, :, None] / L))
    v = bf(rnd(nb, heads, L, d, seed=13))
    ref = _attn_ref(q, k, v).permute(0, 2, 1, 3).reshape(nb * L, heads * d)

    def pack(t):
        p = torch.zeros(nb, L, heads, dpad)
        p[..., :d] = t.permute(0, 2, 1, 3)
        return p.reshape(nb * L, heads * dpad).to(DEV, torch.bfloat16)

    out = o.attention(pack(q), pack(k), pack(v), nb, heads, L, L, d, dpad, 1)
    assert rel(out.cpu(), ref) < 1e-2


@pytest.mark.parametrize("mode,tol", MODES)
@pytest.mark.parametrize("B,Fr,hw,C,heads", [(2, 16, 64, 320, 8), (1, 16, 16, 1280, 8), (2, 8, 9, 640, 8),
                                             (1, 32, 4, 1280, 8), (2, 4, 16, 64, 8), (1, 16, 5, 1280, 8),
                                             (1, 20, 7, 320, 8),
                                             # the instantiated head widths (40 / 80 / 160) with one and two frame tiles,
                                             # odd position counts, fewer heads, and a generic width with a half k-step
                                             (2, 32, 8, 640, 8), (1, 24, 3, 320, 4), (3, 16, 33, 320, 8),
                                             (1, 16, 10, 192, 8), (2, 16, 1024, 320, 8), (1, 12, 6, 1280, 8)])
def test_temporal_attention(mode, tol, B, Fr, hw, C, heads):
    o = ops(mode)
    d = C // heads
    qkv = bf(rnd(B * Fr * hw, 3 * C, seed=1))
    t = qkv.view(B, Fr, hw, 3, heads, d).permute(3, 0, 2, 4, 1, 5)       # [3, B, hw, heads, F, d]
    ref = _attn_ref(t[0], t[1], t[2])                                      # [B, hw, heads, F, d]
    ref = ref.permute(0, 3, 1, 2, 4).reshape(B * Fr * hw, C)
    out = o.temporal_attention(qkv.to(DEV, o.op_dtype), B, Fr, hw, C, heads)
    assert rel(out.cpu(), ref) < (5e-3 if mode == "bf16" else 1e-5)


def test_small_ops():
    o = ops("fp32")
    t = torch.tensor([961.0, 1.0, 500.5])
    half = 160
    freqs = torch.exp(-math.log(10000.0) * torch.arange(half, dtype=torch.float32) / half)
    e = t[:, None] * freqs[None]
    ref = torch.cat([torch.cos(e), torch.sin(e)], -1)
    out = o.timestep_sinusoid(t.to(DEV), 320)
    assert (out.cpu() - ref).abs().max() < 2e-4      # fp32 sin/cos of arguments up to ~1e3
    x, w, b = rnd(3, 320, seed=1), rnd(1280, 320, seed=2, scale=0.05), rnd(1280, seed=3)
    ref = F.silu(F.linear(F.silu(x), w, b))
    out = o.small_linear(x.to(DEV), w.to(DEV), b.to(DEV), silu_in=True, silu_out=True)
    assert rel(out.cpu(), ref) < 1e-5


@pytest.mark.parametrize("mode,tol", MODES)
@pytest.mark.parametrize("B,Fr,h,w", [(2, 3, 16, 12), (1, 2, 5, 7), (2, 16, 32, 32)])
def test_conv_in_out(mode, tol, B, Fr, h, w):
    o = ops(mode)
    s = rnd(B, 4, Fr, h, w, seed=1)
    wi, bi = rnd(320, 4, 3, 3, seed=2, scale=0.2), rnd(320, seed=3)
    x = s.permute(0, 2, 1, 3, 4).reshape(B * Fr, 4, h, w)
    ref = F.conv2d(x, wi, bi, padding=1).permute(0, 2, 3, 1).reshape(B * Fr * h * w, 320)
    out = o.conv_in(s.to(DEV), wi.to(DEV), bi.to(DEV))
    assert rel(out.cpu(), ref) < 1e-5                                     # conv_in is fp32 SIMT in both modes
    act = bf(rnd(B * Fr, h, w, 320, seed=4))
    # bf16 mode: conv_out runs on warp-level MMAs with bf16 operands -> identical (bf16-representable) weights
    wo, bo = bf(rnd(4, 320, 3, 3, seed=5, scale=0.02)), rnd(4, seed=6)
    y = F.conv2d(act.permute(0, 3, 1, 2), wo, bo, padding=1)             # [B*F, 4, h, w]
    ref = y.view(B, Fr, 4, h, w).permute(0, 2, 1, 3, 4)
    out = o.conv_out(act.to(DEV, o.op_dtype).view(-1, 320), B, Fr, h, w, wo.to(DEV), bo.to(DEV))
    assert rel(out.cpu(), ref) < 1e-5                                     # exact products, fp32 accumulation


def test_cfg_step_matches_formula():
    o = ops("fp32")
    n = 4 * 16 * 64 * 64
    u, c, x = rnd(n, seed=1), rnd(n, seed=2), rnd(n, seed=3)
    g, cx, ce = 7.5, 1.0173, -0.2231
    ref = cx * x + ce * (u + g * (c - u))
    mi = torch.empty(2, n, device=DEV)
    out = o.cfg_step(u.to(DEV), c.to(DEV), x.to(DEV), g, cx, ce, in_scale=0.5, model_in=mi)
    assert (out.cpu() - ref).abs().max() < 1e-5
    assert torch.equal(mi[0], mi[1]) and (mi[0].cpu() - 0.5 * ref).abs().max() < 1e-5

"""Temporal-consistency metrics on the device (-m gpu) against the CPU restatement of the
reference's experiments/06_measure_grid_search.py (oracle/metrics.py, oracle/lpips_alex.py)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _video(F, H, W, seed=0):
    """A textured pattern that translates and brightens from frame to frame plus a little noise: non-trivial optical
    flow, warp error, MSE and flicker."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    frames = []
    for f in range(F):
        base = 0.5 + 0.25 * np.sin((xx + 3.0 * f) / 7.0) * np.cos((yy - 2.0 * f) / 9.0) + 0.01 * f * (f % 3)
        img = np.stack([base, 0.9 * base + 0.05, np.roll(base, f, axis=1)], -1) + rng.normal(0, 0.01, (H, W, 3))
        frames.append(np.clip(img, 0, 1))
    return torch.from_numpy((np.stack(frames) * 255).round().astype(np.uint8))


def _close(a, b, tol):
    return abs(a - b) <= tol * max(abs(b), 1e-12)


@pytest.mark.parametrize("F,H,W", [(6, 64, 96), (4, 128, 128), (3, 72, 40)])
def test_video_metrics_match_the_reference_formulas(F, H, W):
    from b200ad import metrics as M
    from oracle import metrics as OM
    u8 = _video(F, H, W)
    frames = u8.permute(0, 3, 1, 2).float() / 255.0                 # what the reference's load_frames returns
    ref = OM.measure_video(frames)
    # the luminance handed to Farneback is produced on the device: identical uint8 images => identical flow
    g = M.gray_u8(u8.cuda()).cpu().numpy()
    g_ref = np.stack([(f.mean(dim=0) * 255).numpy().astype(np.uint8) for f in frames])
    assert (g != g_ref).mean() < 1e-3 and np.abs(g.astype(int) - g_ref.astype(int)).max() <= 1
    out = M.measure_video(u8.cuda())
    for k in ("mean_mse", "std_mse", "mean_psnr", "mean_flow_magnitude", "flow_magnitude_variance", "mean_warp_error",
              "warp_error_variance", "flicker_index"):
        assert _close(out[k], ref[k], 2e-3 if "flow" in k or "warp" in k else 1e-5), (k, out[k], ref[k])
    for po, pr in zip(out["frame_metrics"], ref["frame_metrics"]):
        assert _close(po["mse"], pr["mse"], 1e-5) and _close(po["psnr"], pr["psnr"], 1e-5)
    # with the SAME flow field (the oracle's) the warp error and flow statistics agree tightly
    flows = np.stack([OM.farneback_flow(frames[i], frames[i + 1]) for i in range(F - 1)]).astype(np.float32)
    ws = M.warp_sums(u8.cuda(), torch.from_numpy(flows).cuda()).cpu().numpy()
    for i in range(F - 1):
        we = OM.compute_mse(OM.warp_frame(frames[i], flows[i]), frames[i + 1])
        assert _close(ws[i, 0] / (H * W * 3), we, 1e-4), (i, ws[i, 0] / (H * W * 3), we)
        st = OM.flow_stats(flows[i])
        assert _close(ws[i, 1] / (H * W), st["magnitude_mean"], 1e-5)
    # exact integer sums
    ps = M.pair_sums(u8.cuda()).cpu().numpy()
    a = u8.numpy().astype(np.int64)
    assert ps[0, 0] == ((a[0] - a[1]) ** 2).sum() and ps[0, 1] == np.abs(a[0] - 2 * a[1] + a[2]).sum()
    assert ps[F - 2, 1] == 0


@pytest.mark.parametrize("F,H,W", [(3, 64, 64), (4, 128, 160)])
def test_lpips_alex_matches_oracle(F, H, W):
    from b200ad.lpips import LpipsAlex
    from oracle.lpips_alex import lpips_consecutive, make_oracle_lpips
    oracle = make_oracle_lpips(seed=4)
    prod = LpipsAlex()
    prod.load_state_dict({k: v for k, v in oracle.state_dict().items() if k not in ("shift", "scale")})
    u8 = _video(F, H, W, seed=1)
    ref = lpips_consecutive(oracle, u8.permute(0, 3, 1, 2).float() / 255.0)
    out = prod.cuda()(u8.cuda())
    print("lpips:", out, ref)
    assert len(out) == F - 1 and all(_close(o, r, 1e-4) for o, r in zip(out, ref))
    # the default weights (no lpips checkpoint on this image) are seeded: the same numbers
    dflt = LpipsAlex.from_local_or_random().cuda()(u8.cuda())
    assert all(_close(o, r, 1e-4) for o, r in zip(dflt, ref))


def test_measure_full_size_video_and_cli(tmp_path):
    """16 frames 512x512 (the sweep's video size) end to end incl. LPIPS, and the reference's file layout through
    ``python -m b200ad.metrics``."""
    import json
    import time
    from PIL import Image
    from b200ad import metrics as M
    from b200ad.lpips import LpipsAlex
    u8 = _video(16, 512, 512, seed=2)
    lp = LpipsAlex.from_local_or_random().cuda()
    M.measure_video(u8.cuda(), lpips_fn=lp)                          # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    m = M.measure_video(u8.cuda(), lpips_fn=lp)
    dt = time.perf_counter() - t0
    print(f"measure_video 16f 512^2 incl. LPIPS + Farneback (host): {dt * 1e3:.0f} ms; flicker {m['flicker_index']:.5f}")
    assert m["num_frames"] == 16 and len(m["frame_metrics"]) == 15 and m["mean_lpips"] > 0
    exp = tmp_path / "in" / "portrait_cfg7.5_steps25"
    (exp / "frames").mkdir(parents=True)
    for i in range(4):
        Image.fromarray(u8[i, :64, :64].numpy()).save(exp / "frames" / f"frame_{i:04d}.png")
    (exp / "config.json").write_text(json.dumps({"experiment_id": exp.name, "video_name": "portrait", "guidance_scale": 7.5,
                                                 "num_inference_steps": 25, "phase": "cfg_ablation"}))
    assert M.main(["--input", str(tmp_path / "in"), "--output", str(tmp_path / "out")]) == 0
    rec = json.loads((tmp_path / "out" / f"{exp.name}_metrics.json").read_text())
    summ = json.loads((tmp_path / "out" / "grid_search_results.json").read_text())
    assert rec["num_frames"] == 4 and len(rec["frame_metrics"]) == 3 and summ[0]["experiment_id"] == exp.name
    assert set(summ[0]) == {"experiment_id", "video_name", "guidance_scale", "num_inference_steps", "phase", "mean_mse",
                            "std_mse", "mean_lpips", "std_lpips", "mean_flow_magnitude", "flow_magnitude_variance",
                            "mean_warp_error", "warp_error_variance", "temporal_consistency_score", "flicker_index"}

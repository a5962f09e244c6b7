"""Temporal-consistency metrics of generated videos on the b200ad kernels -- row f4 of SURVEY.md section 8, the
downstream consumer of the sweep (/root/reference/experiments/06_measure_grid_search.py).

``measure_video(frames_u8)`` takes the uint8 frames ``[F, H, W, 3]`` as they sit on the device after the VAE decode
(or as loaded from the PNGs) and returns the reference's ``VideoMetrics`` fields (06:54-80): per frame pair MSE / PSNR /
LPIPS / flow magnitude / warp error, their aggregates, the composite temporal-consistency score and the flicker index.
MSE, flicker and the flow-warped error are fused one-read reductions (csrc/metrics.cu); the dense optical flow is
OpenCV's Farneback on the host, as in the reference (its inputs, the uint8 luminance images, are produced on the
device), the pairs in parallel threads; LPIPS (AlexNet features) runs on the conv kernels (lpips.py) when given.
``python -m b200ad.metrics --input <sweep dir> --output <dir>`` mirrors the reference's command line and file layout
(``<experiment>_metrics.json`` + ``grid_search_results.json``, 06:396-462).
"""
from __future__ import annotations

import argparse
import json
import os
from pathlib import Path
from typing import Callable, List, Optional

import numpy as np
import torch

from . import lib as _lib
from .ops import _empty, _stream


def _farneback(g1: np.ndarray, g2: np.ndarray) -> np.ndarray:
    import cv2
    return cv2.calcOpticalFlowFarneback(g1, g2, None, pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5,
                                        poly_sigma=1.2, flags=0)                 # the reference's parameters, 06:166-177


def gray_u8(frames_u8: torch.Tensor) -> torch.Tensor:
    """uint8 [F,H,W,C] (CUDA) -> uint8 [F,H,W]: the luminance the reference hands to Farneback (06:163-164)."""
    F, H, W, C = frames_u8.shape
    out = _empty((F, H, W), frames_u8.device, torch.uint8)
    _lib.check(_lib.load().b200ad_video_gray_u8(frames_u8.data_ptr(), F * H * W, C, out.data_ptr(), _stream()),
               "b200ad_video_gray_u8")
    return out


def pair_sums(frames_u8: torch.Tensor) -> torch.Tensor:
    """-> float64 [F-1, 2]: sum (I_i - I_{i+1})^2 and sum |I_i - 2 I_{i+1} + I_{i+2}| per pair, uint8 units, exact."""
    L = _lib.load()
    F = frames_u8.shape[0]
    n = frames_u8[0].numel()
    ws = _empty((L.b200ad_metric_blocks() * 3 * (F - 1),), frames_u8.device, torch.float64)
    out = _empty((F - 1, 2), frames_u8.device, torch.float64)
    _lib.check(L.b200ad_video_pair_sums(frames_u8.data_ptr(), F, n, ws.data_ptr(), out.data_ptr(), _stream()),
               "b200ad_video_pair_sums")
    return out


def warp_sums(frames_u8: torch.Tensor, flow: torch.Tensor) -> torch.Tensor:
    """flow fp32 [F-1,H,W,2] -> float64 [F-1, 3]: sum (warp(I_i) - I_{i+1})^2 in [0,1] units, sum |flow|, sum |flow|^2."""
    L = _lib.load()
    F, H, W, C = frames_u8.shape
    assert flow.shape == (F - 1, H, W, 2) and flow.dtype == torch.float32 and flow.is_contiguous()
    ws = _empty((L.b200ad_metric_blocks() * 3 * (F - 1),), frames_u8.device, torch.float64)
    out = _empty((F - 1, 3), frames_u8.device, torch.float64)
    _lib.check(L.b200ad_video_warp_error(frames_u8.data_ptr(), F, H, W, C, flow.data_ptr(), ws.data_ptr(), out.data_ptr(),
                                         _stream()), "b200ad_video_warp_error")
    return out


@torch.no_grad()
def measure_video(frames_u8: torch.Tensor, lpips_fn: Optional[Callable] = None, flow_fn: Callable = _farneback,
                  flow_threads: int = 8) -> dict:
    """frames_u8: uint8 [F, H, W, 3] CUDA tensor.  Returns the fields of the reference's VideoMetrics (06:54-80)."""
    if not frames_u8.is_cuda or frames_u8.dtype != torch.uint8 or frames_u8.dim() != 4:
        raise RuntimeError("measure_video needs a uint8 [F,H,W,C] CUDA tensor (no CPU fallback exists)")
    frames_u8 = frames_u8.contiguous()
    F, H, W, C = frames_u8.shape
    n = H * W * C
    ps = pair_sums(frames_u8)
    gray = gray_u8(frames_u8).cpu().numpy()                                    # host work: Farneback only
    if flow_threads > 1 and F > 2:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=flow_threads) as pool:
            flows = list(pool.map(lambda i: flow_fn(gray[i], gray[i + 1]), range(F - 1)))
    else:
        flows = [flow_fn(gray[i], gray[i + 1]) for i in range(F - 1)]
    flow = torch.from_numpy(np.ascontiguousarray(np.stack(flows).astype(np.float32))).cuda()
    wsum = warp_sums(frames_u8, flow).cpu().numpy()
    ps = ps.cpu().numpy()
    lp = lpips_fn(frames_u8) if lpips_fn is not None else [0.0] * (F - 1)      # [F-1] LPIPS of consecutive frames
    npix = H * W
    pairs = []
    for i in range(F - 1):
        mse = float(ps[i, 0] / (255.0 * 255.0) / n)
        mean_mag = float(wsum[i, 1] / npix)
        var_mag = max(float(wsum[i, 2] / npix) - mean_mag * mean_mag, 0.0)
        pairs.append({"frame_idx": i, "mse": mse, "psnr": 100.0 if mse < 1e-10 else float(10 * np.log10(1.0 / mse)),
                      "lpips": float(lp[i]), "flow_magnitude_mean": mean_mag, "flow_magnitude_std": var_mag ** 0.5,
                      "warp_error": float(wsum[i, 0] / n)})
    mse_v = [p["mse"] for p in pairs]
    lp_v = [p["lpips"] for p in pairs]
    fm = [p["flow_magnitude_mean"] for p in pairs]
    we = [p["warp_error"] for p in pairs]
    flicker = float(np.mean(ps[:F - 2, 1] / 255.0 / n)) if F >= 3 else 0.0
    score = float(np.var(mse_v)) * 1000 + float(np.mean(mse_v)) * 100 + float(np.mean(lp_v)) * 50 + float(np.var(lp_v)) * 500
    return {"num_frames": F, "frame_metrics": pairs,
            "mean_mse": float(np.mean(mse_v)), "std_mse": float(np.std(mse_v)),
            "mean_psnr": float(np.mean([p["psnr"] for p in pairs])),
            "mean_lpips": float(np.mean(lp_v)), "std_lpips": float(np.std(lp_v)),
            "mean_flow_magnitude": float(np.mean(fm)), "flow_magnitude_variance": float(np.var(fm)),
            "mean_warp_error": float(np.mean(we)), "warp_error_variance": float(np.var(we)),
            "temporal_consistency_score": score, "flicker_index": flicker}


def load_frames_u8(frame_dir: Path) -> torch.Tensor:
    """frames/*.png (or *.jpg) -> uint8 [F,H,W,3] on the device (the reference's load_frames, 06:79-95, kept in uint8)."""
    from PIL import Image
    files = sorted(Path(frame_dir).glob("*.png")) or sorted(Path(frame_dir).glob("*.jpg"))
    if not files:
        raise ValueError(f"No frames found in {frame_dir}")
    arr = np.stack([np.asarray(Image.open(f).convert("RGB")) for f in files])
    return torch.from_numpy(arr).cuda()


_SUMMARY_KEYS = ("mean_mse", "std_mse", "mean_lpips", "std_lpips", "mean_flow_magnitude", "flow_magnitude_variance",
                 "mean_warp_error", "warp_error_variance", "temporal_consistency_score", "flicker_index")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="Measure grid search experiments (b200ad kernels)")
    ap.add_argument("--input", default="outputs/05_grid_search")
    ap.add_argument("--output", default="outputs/06_grid_search_metrics")
    ap.add_argument("--filter", default=None)
    ap.add_argument("--no-lpips", action="store_true")
    a = ap.parse_args(argv)
    inp, out = Path(a.input), Path(a.output)
    out.mkdir(parents=True, exist_ok=True)
    if not inp.exists():
        print(f"Error: {inp} does not exist")
        return 1
    dirs = sorted(d for d in inp.iterdir() if d.is_dir() and (d / "frames").exists() and (not a.filter or a.filter in d.name))
    print(f"Found {len(dirs)} experiments to measure")
    lp = None
    if not a.no_lpips:
        from .lpips import LpipsAlex
        lp = LpipsAlex.from_local_or_random().cuda()
    summary = []
    for i, d in enumerate(dirs):
        print(f"\n[{i + 1}/{len(dirs)}] {d.name}")
        with open(d / "config.json") as f:
            cfg = json.load(f)
        m = measure_video(load_frames_u8(d / "frames"), lpips_fn=lp)
        head = {"video_name": cfg["video_name"], "experiment_id": cfg["experiment_id"], "num_frames": m["num_frames"],
                "guidance_scale": cfg.get("guidance_scale", 0), "num_inference_steps": cfg.get("num_inference_steps", 0),
                "phase": cfg.get("phase", "unknown")}
        rec = {**head, **{k: m[k] for k in ("mean_mse", "std_mse", "mean_psnr") + _SUMMARY_KEYS[2:]},
               "frame_metrics": m["frame_metrics"]}
        with open(out / f"{cfg['experiment_id']}_metrics.json", "w", encoding="utf-8") as f:
            json.dump(rec, f, indent=2)
        summary.append({**{k: head[k] for k in ("experiment_id", "video_name", "guidance_scale", "num_inference_steps", "phase")},
                        **{k: m[k] for k in _SUMMARY_KEYS}})
    with open(out / "grid_search_results.json", "w", encoding="utf-8") as f:
        json.dump(summary, f, indent=2)
    print(f"\nResults saved: {out / 'grid_search_results.json'}\nTotal experiments measured: {len(summary)}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())

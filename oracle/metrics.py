"""Oracle: CPU restatement of the reference's temporal-consistency metrics
(/root/reference/experiments/06_measure_grid_search.py) -- SURVEY.md section 8 row f4.

TEST INFRASTRUCTURE ONLY.  The reference module itself cannot be imported here (it imports ``lpips`` at line 19, which
is not installed), so its functions are restated with the SAME torch / numpy / OpenCV calls, each citing its lines:
frames are ``[F, C, H, W]`` float tensors in [0, 1] as ``load_frames`` (06:79-95) returns them.  LPIPS is restated in
oracle/lpips_alex.py (weights absent offline: parity unpinned for that one metric).
"""
from __future__ import annotations

from typing import Callable, List, Optional

import numpy as np
import torch
import torch.nn.functional as F


def compute_mse(a: torch.Tensor, b: torch.Tensor) -> float:                 # 06:209-211
    return F.mse_loss(a, b).item()


def compute_psnr(mse: float) -> float:                                      # 06:214-218
    if mse < 1e-10:
        return 100.0
    return 10 * np.log10(1.0 / mse)


def compute_flicker_index(frames: torch.Tensor) -> float:                   # 06:221-235
    if len(frames) < 3:
        return 0.0
    scores = []
    for t in range(len(frames) - 2):
        second = frames[t] - 2 * frames[t + 1] + frames[t + 2]
        scores.append(second.abs().mean().item())
    return float(np.mean(scores))


def farneback_flow(f1: torch.Tensor, f2: torch.Tensor) -> np.ndarray:       # 06:152-180
    import cv2
    g1 = (f1.mean(dim=0) * 255).numpy().astype(np.uint8)
    g2 = (f2.mean(dim=0) * 255).numpy().astype(np.uint8)
    return cv2.calcOpticalFlowFarneback(g1, g2, None, pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5,
                                        poly_sigma=1.2, flags=0)


def flow_stats(flow: np.ndarray) -> dict:                                   # 06:182-199
    mag = np.sqrt(flow[..., 0] ** 2 + flow[..., 1] ** 2)
    return {"magnitude_mean": float(mag.mean()), "magnitude_std": float(mag.std())}


def warp_frame(frame: torch.Tensor, flow: np.ndarray) -> torch.Tensor:      # 06:258-284
    C, H, W = frame.shape
    gy, gx = np.mgrid[0:H, 0:W].astype(np.float32)
    sx = 2 * (gx + flow[..., 0]) / (W - 1) - 1
    sy = 2 * (gy + flow[..., 1]) / (H - 1) - 1
    grid = torch.stack([torch.from_numpy(sx), torch.from_numpy(sy)], dim=-1).unsqueeze(0)
    return F.grid_sample(frame.unsqueeze(0), grid, mode="bilinear", padding_mode="border", align_corners=True).squeeze(0)


def consistency_score(mse: List[float], lpips: List[float]) -> float:       # 06:238-251
    return float(np.var(mse)) * 1000 + float(np.mean(mse)) * 100 + float(np.mean(lpips)) * 50 + float(np.var(lpips)) * 500


def measure_video(frames: torch.Tensor, lpips_fn: Optional[Callable] = None,
                  flow_fn: Callable = farneback_flow) -> dict:               # 06:290-386
    n = len(frames)
    pairs = []
    for i in range(n - 1):
        f1, f2 = frames[i], frames[i + 1]
        mse = compute_mse(f1, f2)
        flow = flow_fn(f1, f2)
        st = flow_stats(flow)
        pairs.append({"frame_idx": i, "mse": mse, "psnr": compute_psnr(mse),
                      "lpips": float(lpips_fn(f1, f2)) if lpips_fn is not None else 0.0,
                      "flow_magnitude_mean": st["magnitude_mean"], "flow_magnitude_std": st["magnitude_std"],
                      "warp_error": compute_mse(warp_frame(f1, flow), f2)})
    mse_v = [p["mse"] for p in pairs]
    lp_v = [p["lpips"] for p in pairs]
    fm = [p["flow_magnitude_mean"] for p in pairs]
    we = [p["warp_error"] for p in pairs]
    return {"num_frames": n, "frame_metrics": pairs,
            "mean_mse": float(np.mean(mse_v)), "std_mse": float(np.std(mse_v)),
            "mean_psnr": float(np.mean([p["psnr"] for p in pairs])),
            "mean_lpips": float(np.mean(lp_v)), "std_lpips": float(np.std(lp_v)),
            "mean_flow_magnitude": float(np.mean(fm)), "flow_magnitude_variance": float(np.var(fm)),
            "mean_warp_error": float(np.mean(we)), "warp_error_variance": float(np.var(we)),
            "temporal_consistency_score": consistency_score(mse_v, lp_v),
            "flicker_index": compute_flicker_index(frames)}

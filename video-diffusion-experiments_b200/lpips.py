"""LPIPS (AlexNet) between consecutive frames on the b200ad kernels -- part of row f4: replaces
``lpips.LPIPS(net='alex')`` of /root/reference/experiments/06_measure_grid_search.py:101-131.

All frames of a video go through the AlexNet feature stack ONCE (the reference evaluates every interior frame twice,
once per pair): b200ad_lpips_input (uint8 -> [-1,1] -> ScalingLayer), five b200ad_conv2d_direct layers with fused ReLU
(11x11 stride 4, 5x5, three 3x3), two b200ad_maxpool2d, and per tapped layer one b200ad_lpips_layer reduction
(unit-normalise, squared difference, lin weights, spatial mean) over the F-1 consecutive pairs.  State-dict keys follow the
lpips package (``net.sliceK.N.{weight,bias}``, ``linK.model.1.weight``) so its checkpoint loads when present; without it
(this image: no package, no weights) the weights are seeded random and the log says so.
"""
from __future__ import annotations

import os
import sys

import torch
import torch.nn as nn

from . import lib as _lib
from .ops import _empty, _stream

CHANNELS = (64, 192, 384, 256, 256)
ALEX_CONVS = ((1, 0, 3, 64, 11, 4, 2, False), (2, 3, 64, 192, 5, 1, 2, True), (3, 6, 192, 384, 3, 1, 1, True),
              (4, 8, 384, 256, 3, 1, 1, False), (5, 10, 256, 256, 3, 1, 1, False))


class _Lin(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.model = nn.Sequential(nn.Dropout(), nn.Conv2d(c, 1, 1, bias=False))


class LpipsAlex(nn.Module):
    def __init__(self):
        super().__init__()
        self.net = nn.Module()
        for sl, idx, cin, cout, k, s, p, _ in ALEX_CONVS:
            seq = nn.Module()
            setattr(seq, str(idx), nn.Conv2d(cin, cout, k, stride=s, padding=p))
            setattr(self.net, f"slice{sl}", seq)
        for i, c in enumerate(CHANNELS):
            setattr(self, f"lin{i}", _Lin(c))
        self.eval()
        self._packed = None
        self.register_load_state_dict_post_hook(lambda m, keys: setattr(m, "_packed", None))

    def _apply(self, fn, *a, **kw):
        r = super()._apply(fn, *a, **kw)
        self._packed = None
        return r

    @classmethod
    def from_local_or_random(cls, path: str = None, seed: int = 4):
        m = cls()
        path = path or os.environ.get("B200AD_LPIPS_WEIGHTS", "")
        if path and os.path.isfile(path):
            sd = torch.load(path, map_location="cpu", weights_only=True)
            sd = {k: v for k, v in sd.items() if not k.startswith("scaling_layer")}
            m.load_state_dict(sd, strict=False)
        else:
            print("[b200ad] LPIPS: no local weights (lpips package / checkpoint absent) -> seeded random AlexNet + lin "
                  "weights; values are comparable between runs of this build only", file=sys.stderr, flush=True)
            state = torch.random.get_rng_state()
            torch.manual_seed(seed)
            ref = cls()
            torch.random.set_rng_state(state)
            with torch.no_grad():
                for i in range(5):
                    getattr(ref, f"lin{i}").model[1].weight.abs_()
            m.load_state_dict(ref.state_dict())
        return m

    def _pack(self):
        if self._packed is None:
            convs = []
            for sl, idx, cin, cout, k, s, p, pool in ALEX_CONVS:
                conv = getattr(getattr(self.net, f"slice{sl}"), str(idx))
                w = conv.weight.detach().float().permute(0, 2, 3, 1).contiguous().cuda()       # [cout, kh, kw, cin]
                convs.append((w, conv.bias.detach().float().contiguous().cuda(), cin, cout, k, s, p, pool))
            lins = [getattr(self, f"lin{i}").model[1].weight.detach().float().reshape(-1).contiguous().cuda()
                    for i in range(5)]
            self._packed = (convs, lins)
        return self._packed

    @torch.no_grad()
    def forward(self, frames_u8: torch.Tensor) -> list:
        """uint8 [F, H, W, 3] CUDA tensor -> [F-1] LPIPS distances of consecutive frames (python floats)."""
        if not frames_u8.is_cuda or frames_u8.dtype != torch.uint8:
            raise RuntimeError("LpipsAlex needs a uint8 [F,H,W,3] CUDA tensor (no CPU fallback exists)")
        L = _lib.load()
        convs, lins = self._pack()
        frames_u8 = frames_u8.contiguous()
        N, H, W, C = frames_u8.shape
        x = _empty((N, H, W, 3), frames_u8.device, torch.float32)
        _lib.check(L.b200ad_lpips_input(frames_u8.data_ptr(), frames_u8.numel(), x.data_ptr(), _stream()), "b200ad_lpips_input")
        per_layer = []
        ws = _empty((L.b200ad_metric_blocks() * (N - 1),), frames_u8.device, torch.float64)
        h, w = H, W
        for li, (wt, b, cin, cout, k, s, p, pool) in enumerate(convs):
            if pool:
                oh, ow = (h - 3) // 2 + 1, (w - 3) // 2 + 1
                y = _empty((N, oh, ow, cin), x.device, torch.float32)
                _lib.check(L.b200ad_maxpool2d(x.data_ptr(), N, h, w, cin, 3, 2, y.data_ptr(), _stream()), "b200ad_maxpool2d")
                x, h, w = y, oh, ow
            oh, ow = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
            y = _empty((N, oh, ow, cout), x.device, torch.float32)
            _lib.check(L.b200ad_conv2d_direct(x.data_ptr(), N, h, w, cin, wt.data_ptr(), b.data_ptr(), cout, k, k, s, p, 1,
                                              y.data_ptr(), _stream()), "b200ad_conv2d_direct")
            x, h, w = y, oh, ow
            d = _empty((N - 1,), x.device, torch.float64)
            _lib.check(L.b200ad_lpips_layer(x.data_ptr(), N, h * w, cout, lins[li].data_ptr(), ws.data_ptr(), d.data_ptr(),
                                            _stream()), "b200ad_lpips_layer")
            per_layer.append((d, h * w))
        host = [(d.cpu().tolist(), hw) for d, hw in per_layer]                 # five layers: spatial mean, then the sum
        return [sum(v[i] / hw for v, hw in host) for i in range(N - 1)]

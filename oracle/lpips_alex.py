"""Oracle: CPU restatement of LPIPS v0.1 with the AlexNet backbone, the perceptual metric the reference computes
between consecutive frames (/root/reference/experiments/06_measure_grid_search.py:101-131: ``lpips.LPIPS(net='alex')``,
inputs mapped from [0,1] to [-1,1]).

TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED for this metric: neither the ``lpips`` package nor its weights exist on this
image; the restatement follows the published module (richzhang/PerceptualSimilarity, lpips/lpips.py + pretrained_networks
.py): ScalingLayer (shift [-.030,-.088,-.188], scale [.458,.448,.450]) -> torchvision AlexNet ``features`` tapped after
each of the five ReLUs (64, 192, 384, 256, 256 channels) -> channel-wise unit normalisation (x / (|x| + 1e-10)) ->
squared difference -> non-negative 1x1 ``lin`` weights -> spatial mean -> sum over the five layers.  The AlexNet layer
geometry is cross-checked against torchvision.models.alexnet (importable) in tests/test_oracle_cpu.py; weights are
random (seeded), state-dict keys follow the lpips package (``net.sliceK.N.weight``, ``linK.model.1.weight``).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

CHANNELS = (64, 192, 384, 256, 256)
# (slice, index inside torchvision's alexnet.features, cin, cout, kernel, stride, padding, maxpool before the conv)
ALEX_CONVS = ((1, 0, 3, 64, 11, 4, 2, False), (2, 3, 64, 192, 5, 1, 2, True), (3, 6, 192, 384, 3, 1, 1, True),
              (4, 8, 384, 256, 3, 1, 1, False), (5, 10, 256, 256, 3, 1, 1, False))


class _Lin(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.model = nn.Sequential(nn.Dropout(), nn.Conv2d(c, 1, 1, bias=False))


class OracleLpipsAlex(nn.Module):
    def __init__(self):
        super().__init__()
        self.register_buffer("shift", torch.tensor([-.030, -.088, -.188])[None, :, None, None])
        self.register_buffer("scale", torch.tensor([.458, .448, .450])[None, :, None, None])
        self.net = nn.Module()
        for sl, idx, cin, cout, k, s, p, _ in ALEX_CONVS:
            seq = nn.Module()
            setattr(seq, str(idx), nn.Conv2d(cin, cout, k, stride=s, padding=p))
            setattr(self.net, f"slice{sl}", seq)
        for i, c in enumerate(CHANNELS):
            setattr(self, f"lin{i}", _Lin(c))

    def features(self, x):
        feats = []
        for sl, idx, *_rest, pool in ALEX_CONVS:
            if pool:
                x = F.max_pool2d(x, kernel_size=3, stride=2)
            conv = getattr(getattr(self.net, f"slice{sl}"), str(idx))
            x = F.relu(conv(x))
            feats.append(x)
        return feats

    @torch.no_grad()
    def forward(self, in0: torch.Tensor, in1: torch.Tensor) -> torch.Tensor:
        """in0, in1: [N, 3, H, W] in [-1, 1] -> [N] distances."""
        f0 = self.features((in0 - self.shift) / self.scale)
        f1 = self.features((in1 - self.shift) / self.scale)
        total = 0
        for i, (a, b) in enumerate(zip(f0, f1)):
            na = a / (a.pow(2).sum(1, keepdim=True).sqrt() + 1e-10)
            nb = b / (b.pow(2).sum(1, keepdim=True).sqrt() + 1e-10)
            w = getattr(self, f"lin{i}").model[1].weight
            total = total + ((na - nb) ** 2 * w).sum(1, keepdim=True).mean((2, 3), keepdim=True)
        return total.flatten()


def make_oracle_lpips(seed: int = 4) -> OracleLpipsAlex:
    state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    m = OracleLpipsAlex().eval()
    with torch.no_grad():
        for i in range(5):
            getattr(m, f"lin{i}").model[1].weight.abs_()          # the published lin weights are non-negative
    torch.random.set_rng_state(state)
    for p in m.parameters():
        p.requires_grad_(False)
    return m


def lpips_consecutive(m: OracleLpipsAlex, frames: torch.Tensor) -> list:
    """frames [F, 3, H, W] in [0, 1] -> [F-1] LPIPS of consecutive frames, as the reference calls it (06:113-127)."""
    x = frames * 2 - 1
    return [float(m(x[i:i + 1], x[i + 1:i + 2])) for i in range(len(frames) - 1)]

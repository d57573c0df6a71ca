"""Feature-space (perceptual) loss.

Parity target: ``loss = feat_loss`` from the absent ``PyTorchPercept``
(Stoke-DDP.py:35,224): a plain callable ``feat_loss(output, target) -> scalar`` built on
VGG feature maps.  There are no pretrained weights on the box (no network), so the
feature extractor is a frozen, deterministically initialised VGG-16 front end; pass
``state_dict`` to load real ImageNet weights when available.  The loss is
``pixel_weight * L1(out, tgt) + sum_k w_k * MSE(phi_k(out), phi_k(tgt))``.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops.layers import Conv2d

__all__ = ["VGGFeatures", "PerceptualLoss", "feat_loss"]

_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512]
_TAPS = (3, 8, 15)  # relu1_2, relu2_2, relu3_3 (indices in torchvision's vgg16.features)


class VGGFeatures(nn.Module):
    def __init__(self, taps: Sequence[int] = _TAPS, seed: int = 0):
        super().__init__()
        layers: List[nn.Module] = []
        c = 3
        for v in _CFG:
            if v == "M":
                layers.append(nn.MaxPool2d(2, 2))
            else:
                layers += [Conv2d(c, int(v), 3, padding=1), nn.ReLU(inplace=False)]
                c = int(v)
        self.features = nn.Sequential(*layers[: max(taps) + 1])
        self.taps = tuple(taps)
        g = torch.Generator().manual_seed(seed)
        for m in self.features:
            if isinstance(m, nn.Conv2d):
                fan_in = m.weight[0].numel()
                m.weight.data = torch.randn(m.weight.shape, generator=g) * (2.0 / fan_in) ** 0.5
                m.bias.data.zero_()
        self.register_buffer("mean", torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1))
        self.register_buffer("std", torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1))
        for p in self.parameters():
            p.requires_grad_(False)

    def forward(self, x: torch.Tensor) -> List[torch.Tensor]:
        x = (x - self.mean.to(x.dtype)) / self.std.to(x.dtype)
        out = []
        for i, m in enumerate(self.features):
            x = m(x)
            if i in self.taps:
                out.append(x)
        return out


class PerceptualLoss(nn.Module):
    def __init__(self, weights: Sequence[float] = (1.0, 1.0, 1.0), pixel_weight: float = 1.0,
                 state_dict: Optional[Dict[str, torch.Tensor]] = None):
        super().__init__()
        self.vgg = VGGFeatures()
        if state_dict is not None:
            self.vgg.features.load_state_dict(state_dict, strict=False)
        self.weights, self.pixel_weight = tuple(weights), pixel_weight
        self.eval()

    def forward(self, output: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        loss = self.pixel_weight * F.l1_loss(output, target)
        fo = self.vgg(output)
        with torch.no_grad():
            ft = self.vgg(target)
        for w, a, b in zip(self.weights, fo, ft):
            loss = loss + w * F.mse_loss(a, b)
        return loss


_CACHE: Dict[str, PerceptualLoss] = {}


def feat_loss(output: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """Module-level callable with the reference's signature (Stoke-DDP.py:224)."""
    key = f"{output.device}"
    mod = _CACHE.get(key)
    if mod is None:
        mod = _CACHE[key] = PerceptualLoss().to(output.device)
    return mod(output, target)

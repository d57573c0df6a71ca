"""Image-quality metrics used by the validation loop.

Parity target: ``metrics.mae(outputs, targets)`` and ``metrics.psnr(outputs, targets)``
(Stoke-DDP.py:120-121); the ``metrics`` module is absent from the reference, so these
implement the standard definitions on [0, 1]-ranged images.  Both return device
scalars (no host sync); ``reduce_mean`` averages a metric across ranks, which the
reference does not do but a multi-GPU validation needs.
"""
from __future__ import annotations

from typing import Any, Optional

import torch
import torch.distributed as dist


@torch.no_grad()
def mae(outputs: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    return (outputs.float() - targets.float()).abs().mean()


@torch.no_grad()
def mse(outputs: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    d = outputs.float() - targets.float()
    return (d * d).mean()


@torch.no_grad()
def psnr(outputs: torch.Tensor, targets: torch.Tensor, data_range: float = 1.0, eps: float = 1e-10) -> torch.Tensor:
    """Mean per-image peak signal-to-noise ratio in dB."""
    d = outputs.float().clamp(0.0, data_range) - targets.float()
    per_image = (d * d).flatten(1).mean(1) if d.dim() > 1 else (d * d).mean().unsqueeze(0)
    return (10.0 * torch.log10(data_range * data_range / (per_image + eps))).mean()


@torch.no_grad()
def accuracy(logits: torch.Tensor, targets: torch.Tensor, topk: int = 1) -> torch.Tensor:
    pred = logits.topk(topk, dim=1).indices
    return (pred == targets.view(-1, 1)).any(1).float().mean()


@torch.no_grad()
def reduce_mean(value: torch.Tensor, group: Optional[Any] = None) -> torch.Tensor:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        value = value.detach().clone()
        dist.all_reduce(value, group=group)
        value /= dist.get_world_size(group)
    return value

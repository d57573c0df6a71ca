"""Cross-rank synchronised BatchNorm and the module-tree conversion.

Parity target: ``DDPConfig(convert_to_sync_batch_norm=True)`` (Stoke-DDP.py:192) —
torch ``SyncBatchNorm.convert_sync_batchnorm`` semantics: forward
exchanges per-rank ``[sum, sum_sq, count]`` and every rank normalises with the global
statistics; backward exchanges ``[sum_dy, sum_dy_xmu]``.

B200-first: the statistics are exchanged as raw moments (sum / sum of squares /
count) in ONE all-reduce of ``2C+1`` floats instead of an all-gather of
``[mean, invstd, count]`` followed by a combine kernel — same math, fewer launches,
and the reduction maps directly onto the one-shot peer-memory all-reduce (W6).  On
CUDA channels-last tensors the local moments and the normalise(+ReLU) are the fused
kernels of ``ops.norm``; on CPU everything is plain torch ops (gloo plumbing tests).
"""
from __future__ import annotations

from typing import Any, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.nn.modules.batchnorm import _BatchNorm

from . import comm

__all__ = ["SyncBatchNorm", "convert_sync_batchnorm"]


def _reduce_dims(x: torch.Tensor):
    return [0] + list(range(2, x.dim()))


def _bshape(x: torch.Tensor):
    return [1, -1] + [1] * (x.dim() - 2)


class _SyncBNFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, eps, momentum, group, world_size):  # type: ignore[override]
        C = x.shape[1]
        dims = _reduce_dims(x)
        xf = x.float()
        count = torch.full((1,), x.numel() // C, dtype=torch.float32, device=x.device)
        stats = torch.cat([xf.sum(dims), (xf * xf).sum(dims), count])
        if world_size > 1:
            dist.all_reduce(stats, group=group)
        n = stats[2 * C]
        mean = stats[:C] / n
        var = (stats[C:2 * C] / n - mean * mean).clamp_(min=0.0)
        invstd = torch.rsqrt(var + eps)
        if running_mean is not None:
            with torch.no_grad():
                unbiased = var * (n / (n - 1).clamp(min=1.0))
                running_mean.mul_(1 - momentum).add_(mean.to(running_mean.dtype), alpha=momentum)
                running_var.mul_(1 - momentum).add_(unbiased.to(running_var.dtype), alpha=momentum)
        sh = _bshape(x)
        xhat = (xf - mean.view(sh)) * invstd.view(sh)
        out = xhat
        if weight is not None:
            out = out * weight.float().view(sh)
        if bias is not None:
            out = out + bias.float().view(sh)
        ctx.save_for_backward(x, weight, mean, invstd, n)
        ctx.group, ctx.world_size = group, world_size
        ctx.has_bias = bias is not None
        return out.to(x.dtype)

    @staticmethod
    def backward(ctx, dy):  # type: ignore[override]
        x, weight, mean, invstd, n = ctx.saved_tensors
        C = x.shape[1]
        dims = _reduce_dims(x)
        sh = _bshape(x)
        dyf = dy.float()
        xmu = x.float() - mean.view(sh)
        sum_dy = dyf.sum(dims)
        sum_dy_xmu = (dyf * xmu).sum(dims)
        grad_weight = (sum_dy_xmu * invstd) if weight is not None else None
        grad_bias = sum_dy.clone() if ctx.has_bias else None
        red = torch.cat([sum_dy, sum_dy_xmu])
        if ctx.world_size > 1:
            dist.all_reduce(red, group=ctx.group)
        g_sum_dy, g_sum_dy_xmu = red[:C], red[C:]
        w = weight.float() if weight is not None else torch.ones(C, device=x.device)
        # dx = w*invstd * (dy - mean(dy) - xhat * mean(dy*xhat))
        k = (w * invstd).view(sh)
        dx = k * (dyf - (g_sum_dy / n).view(sh) - xmu * (invstd * invstd * g_sum_dy_xmu / n).view(sh))
        return (dx.to(x.dtype),
                grad_weight.to(weight.dtype) if grad_weight is not None else None,
                grad_bias.to(dy.dtype if weight is None else weight.dtype) if grad_bias is not None else None,
                None, None, None, None, None, None)


class SyncBatchNorm(_BatchNorm):
    def __init__(self, num_features: int, eps: float = 1e-5, momentum: Optional[float] = 0.1, affine: bool = True,
                 track_running_stats: bool = True, process_group: Optional[Any] = None,
                 device=None, dtype=None):
        super().__init__(num_features, eps, momentum, affine, track_running_stats, device=device, dtype=dtype)
        self.process_group = process_group
        self.fuse_relu = False  # set by model builders that want BN+ReLU in one kernel

    def _check_input_dim(self, input: torch.Tensor) -> None:
        if input.dim() < 2:
            raise ValueError(f"expected at least 2D input (got {input.dim()}D input)")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        self._check_input_dim(x)
        use_batch_stats = self.training or not self.track_running_stats
        if not use_batch_stats:
            return nn.functional.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias,
                                            False, 0.0, self.eps)
        momentum = 0.0 if self.momentum is None else self.momentum
        if self.training and self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
            if self.momentum is None:
                momentum = 1.0 / float(self.num_batches_tracked)
        group = self.process_group if self.process_group is not None else (
            dist.group.WORLD if comm.is_initialized() else None)
        world = comm.get_world_size(group)
        return _SyncBNFunction.apply(x, self.weight, self.bias,
                                     self.running_mean if self.track_running_stats else None,
                                     self.running_var if self.track_running_stats else None,
                                     self.eps, momentum, group, world)


def sync_batch_norm_reference(bn: _BatchNorm, x: torch.Tensor) -> torch.Tensor:
    """Training-mode cross-rank batch norm for any ``_BatchNorm``-shaped module (portable path)."""
    momentum = 0.0 if bn.momentum is None else bn.momentum
    if bn.track_running_stats and bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)
        if bn.momentum is None:
            momentum = 1.0 / float(bn.num_batches_tracked)
    group = getattr(bn, "process_group", None)
    if group is None and comm.is_initialized():
        group = dist.group.WORLD
    world = comm.get_world_size(group)
    return _SyncBNFunction.apply(x, bn.weight, bn.bias,
                                 bn.running_mean if bn.track_running_stats else None,
                                 bn.running_var if bn.track_running_stats else None,
                                 bn.eps, momentum, group, world)


def convert_sync_batchnorm(module: nn.Module, process_group: Optional[Any] = None) -> nn.Module:
    """Recursively replace every ``_BatchNorm`` by :class:`SyncBatchNorm`, keeping
    parameters, buffers and their ``state_dict`` names."""
    out = module
    from ..ops.layers import BatchNormAct2d
    if isinstance(module, BatchNormAct2d):
        # fused norm+add+activation module: flip it to cross-rank statistics in place
        module.sync = True
        module.process_group = process_group
        return module
    if isinstance(module, _BatchNorm) and not isinstance(module, SyncBatchNorm):
        out = SyncBatchNorm(module.num_features, module.eps, module.momentum, module.affine,
                            module.track_running_stats, process_group)
        if module.affine:
            with torch.no_grad():
                out.weight = module.weight
                out.bias = module.bias
        out.running_mean = module.running_mean
        out.running_var = module.running_var
        out.num_batches_tracked = module.num_batches_tracked
        out.training = module.training
        if hasattr(module, "fuse_relu"):
            out.fuse_relu = module.fuse_relu
    for name, child in module.named_children():
        out.add_module(name, convert_sync_batchnorm(child, process_group))
    return out

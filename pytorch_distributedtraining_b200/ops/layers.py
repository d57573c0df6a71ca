"""``nn.Module`` building blocks the model zoo is written against.

They keep stock parameter names / shapes (so state dicts interchange with torchvision
style checkpoints) and choose the implementation per call:

  * CUDA tensors in the supported layout/dtype → the hand-written sm_100a kernels
    (``ops.norm``, ``ops.gemm``, ``ops.conv``) through autograd functions;
  * anything else (CPU, fp64, exotic shapes) → stock torch ops.  On a GPU box a missing
    kernel library is an error, not a silent fallback (``ops.require_library``).

Kernel work-list rows served here: W7 (implicit-GEMM conv), W8 (BN+ReLU), K17 (linear).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import dispatch


class Conv2d(nn.Conv2d):
    """Convolution.  On CUDA channels-last fp16/bf16 the forward/dgrad/wgrad go through
    ``ops.conv`` (tcgen05 implicit GEMM) when the shape is supported; otherwise cuDNN."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and dispatch.conv_supported(self, x):
            from . import conv as _conv
            return _conv.conv2d(x, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups)
        return super().forward(x)


class Linear(nn.Linear):
    """Linear layer; tcgen05 GEMM with fused bias (+activation) epilogue on CUDA."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, act: Optional[str] = None,
                 device=None, dtype=None):
        super().__init__(in_features, out_features, bias, device=device, dtype=dtype)
        self.act = act

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and dispatch.linear_supported(self, x):
            from . import gemm as _gemm
            return _gemm.linear(x, self.weight, self.bias, act=self.act)
        y = F.linear(x, self.weight, self.bias)
        if self.act == "gelu":
            y = F.gelu(y)
        elif self.act == "relu":
            y = F.relu(y)
        return y


class BatchNormAct2d(nn.BatchNorm2d):
    """BatchNorm2d with an optional fused residual add and activation:
    ``y = act(bn(x) + residual)``.  Same parameters/buffers as ``nn.BatchNorm2d``."""

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: float = 0.1, affine: bool = True,
                 track_running_stats: bool = True, act: Optional[str] = "relu", device=None, dtype=None):
        super().__init__(num_features, eps, momentum, affine, track_running_stats, device=device, dtype=dtype)
        self.act = act
        self.fuse_relu = act == "relu"
        self.process_group = None  # set by convert_sync_batchnorm → cross-rank statistics
        self.sync = False

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:  # type: ignore[override]
        if x.is_cuda and dispatch.bn_supported(self, x):
            from . import norm as _norm
            return _norm.batch_norm_act(self, x, residual)
        if self.sync and self.training:
            from ..parallel.sync_bn import sync_batch_norm_reference
            y = sync_batch_norm_reference(self, x)
        else:
            y = super().forward(x)
        if residual is not None:
            y = y + residual
        if self.act == "relu":
            y = F.relu(y)
        return y


class LayerNorm(nn.LayerNorm):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and dispatch.ln_supported(self, x):
            from . import norm as _norm
            return _norm.layer_norm(x, self.weight, self.bias, self.eps)
        return super().forward(x)

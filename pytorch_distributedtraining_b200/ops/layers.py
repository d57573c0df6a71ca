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

    # set by ``pair_conv_bn``: a BatchNormAct2d consumes this layer's output directly, so in training
    # the convolution's epilogue also emits the batch statistics that layer needs
    emit_bn_stats: bool = False
    # "physical" channel counts (set by a model that keeps its activations zero-padded to a tensor-core friendly width,
    # e.g. SwinIR-S 60 -> 64): the input then HAS phys_in channels (extras are zeros) and the output is produced with
    # phys_out channels (extras are zeros) — only the small filters are padded, never an activation
    phys_in: Optional[int] = None
    phys_out: Optional[int] = None

    def _forward_phys(self, x: torch.Tensor) -> torch.Tensor:
        from .fused import padded_param
        cin, cout = self.in_channels, self.out_channels
        pi, po = self.phys_in or cin, self.phys_out or cout
        cache = self.__dict__.setdefault("_phys_cache", {"w": {}, "b": {}})
        lp = torch.get_autocast_dtype("cuda") if (x.is_cuda and torch.is_autocast_enabled()) else None
        w = padded_param(self.weight, (po, pi) + tuple(self.weight.shape[2:]), cache["w"], lp)
        b = None if self.bias is None else padded_param(self.bias, (po,), cache["b"])
        if x.is_cuda and dispatch.conv_phys_supported(self, x, pi, po):
            from . import conv as _conv
            return _conv.conv2d(x, w, b, self.stride, self.padding, self.dilation, self.groups)
        return F.conv2d(x, w, b, self.stride, self.padding, self.dilation, self.groups)

    def forward(self, x: torch.Tensor, carry: bool = False):
        if (self.phys_in or self.phys_out) and not carry:
            return self._forward_phys(x)
        if x.is_cuda and dispatch.conv_supported(self, x):
            from . import conv as _conv
            emit = self.emit_bn_stats and self.training and torch.is_grad_enabled() and dispatch.enabled("bn") \
                and dispatch.enabled("convstats")
            return _conv.conv2d(x, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups,
                                emit_bn_stats=emit, carry=carry)
        if x.is_cuda and not carry and dispatch.conv_padded_supported(self, x):
            return self._forward_padded(x)
        y = super().forward(x)
        return (y, x) if carry else y

    def _forward_padded(self, x: torch.Tensor) -> torch.Tensor:
        """Channel counts just outside the kernel envelope (60 -> 64): zero-pad input channels and filters, run the
        implicit-GEMM kernels, slice the output channels.  The pads / slice are ordinary autograd ops."""
        from . import conv as _conv
        cin, cout = self.in_channels, self.out_channels
        first_layer = cin <= 4                       # RGB stems keep their channels: only Cout is padded (to 8)
        cin_p = cin if first_layer else (cin + 63) // 64 * 64
        cout_p = (cout + 63) // 64 * 64 if (x.requires_grad and not first_layer) else (cout + 7) // 8 * 8
        dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else x.dtype
        xp = x.to(dt)
        if cin_p != cin:
            xp = torch.cat([xp, xp.new_zeros((x.shape[0], cin_p - cin, x.shape[2], x.shape[3]))], dim=1)
        xp = xp.contiguous(memory_format=torch.channels_last)
        w = F.pad(self.weight, (0, 0, 0, 0, 0, cin_p - cin, 0, cout_p - cout))
        b = None if self.bias is None else F.pad(self.bias, (0, cout_p - cout))
        y = _conv.conv2d(xp, w, b, self.stride, self.padding, self.dilation, self.groups)
        return y[:, :cout] if cout_p != cout else y

    def forward_carry(self, x: torch.Tensor):
        """``(conv(x), x)``: use the second value for the branch that bypasses this layer (residual identity,
        downsample).  On the kernel path its gradient is routed back through this layer's backward and the
        input gradient is accumulated into it in the GEMM epilogue, which removes autograd's add kernel."""
        return self.forward(x, carry=True)


class Linear(nn.Linear):
    """Linear layer; tcgen05 GEMM with fused bias (+activation) epilogue on CUDA."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, act: Optional[str] = None,
                 device=None, dtype=None):
        super().__init__(in_features, out_features, bias, device=device, dtype=dtype)
        self.act = act
        # physical widths, as for Conv2d: the input has phys_in features (extras are zeros), the output phys_out
        self.phys_in: Optional[int] = None
        self.phys_out: Optional[int] = None

    def padded_params(self):
        """Weight / bias zero-padded to the physical widths (autograd slices the gradients back)."""
        from .fused import padded_param
        k, n = self.in_features, self.out_features
        kp, np_ = self.phys_in or k, self.phys_out or n
        if kp == k and np_ == n:
            return self.weight, self.bias
        cache = self.__dict__.setdefault("_phys_cache", {"w": {}, "b": {}})
        lp = torch.get_autocast_dtype("cuda") if (self.weight.is_cuda and torch.is_autocast_enabled()) else None
        w = padded_param(self.weight, (np_, kp), cache["w"], lp)
        b = None if self.bias is None else padded_param(self.bias, (np_,), cache["b"])
        return w, b

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``act(x W^T + b) (+ residual)``; on the kernel path the residual add is part of the GEMM epilogue."""
        if self.phys_in or self.phys_out:
            w, b = self.padded_params()
            if x.is_cuda and dispatch.linear_phys_supported(w, x):
                from . import gemm as _gemm
                return _gemm.linear(x, w, b, act=self.act, residual=residual)
            y = F.linear(x, w.to(x.dtype) if w.dtype != x.dtype and not torch.is_autocast_enabled() else w,
                         None if b is None else (b.to(x.dtype) if b.dtype != x.dtype and not torch.is_autocast_enabled() else b))
            if self.act == "gelu":
                y = F.gelu(y)
            elif self.act == "relu":
                y = F.relu(y)
            return y if residual is None else y + residual
        if x.is_cuda and dispatch.linear_supported(self, x):
            from . import gemm as _gemm
            return _gemm.linear(x, self.weight, self.bias, act=self.act, residual=residual)
        if x.is_cuda and dispatch.linear_padded_supported(self, x):
            # widths that are not multiples of 8 (SwinIR-S: 60 / 180 / 120): zero-pad K and N to the next multiple of 8,
            # run the tcgen05 GEMM, slice — the pads and the slice are ordinary autograd ops
            from . import gemm as _gemm
            k, n = self.in_features, self.out_features
            kp, np_ = (k + 7) // 8 * 8, (n + 7) // 8 * 8
            xp = F.pad(x, (0, kp - k)) if kp != k else x
            w = F.pad(self.weight, (0, kp - k, 0, np_ - n)) if (kp != k or np_ != n) else self.weight
            b = None if self.bias is None else (F.pad(self.bias, (0, np_ - n)) if np_ != n else self.bias)
            y = _gemm.linear(xp, w, b, act=self.act)
            if np_ != n:
                y = y[..., :n]
            return y if residual is None else y + residual
        y = F.linear(x, self.weight, self.bias)
        if self.act == "gelu":
            y = F.gelu(y)
        elif self.act == "relu":
            y = F.relu(y)
        return y if residual is None else y + residual


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


def pair_conv_bn(model: nn.Module) -> int:
    """Mark every ``Conv2d`` that is registered immediately before a ``BatchNormAct2d`` in the same parent
    (conv1/bn1, Sequential(conv, bn) ...) so that it emits the batch statistics from its epilogue.
    The pairing is only a hint: the BatchNorm uses the vector only when it finds it on its own input
    tensor, so a wrong guess costs a little epilogue work and nothing else.  Returns the pair count."""
    pairs = 0
    for parent in model.modules():
        prev = None
        for child in parent.children():
            if isinstance(child, BatchNormAct2d) and isinstance(prev, Conv2d) and prev.out_channels == child.num_features:
                prev.emit_bn_stats = True
                pairs += 1
            prev = child
    return pairs


class LayerNorm(nn.LayerNorm):
    # physical row width (see Conv2d.phys_in): rows of the input are `phys` wide, the first normalized_shape[0] elements
    # are normalised and the rest of the output row is zeros
    phys: Optional[int] = None

    def _forward_phys(self, x: torch.Tensor) -> torch.Tensor:
        d = self.normalized_shape[0]
        if x.is_cuda and dispatch.ln_phys_supported(self, x):
            from . import norm as _norm
            return _norm.layer_norm(x, self.weight, self.bias, self.eps, logical=d)
        return F.pad(F.layer_norm(x[..., :d], (d,), self.weight, self.bias, self.eps), (0, x.shape[-1] - d))

    def forward(self, x: torch.Tensor, carry: bool = False):
        if self.phys and self.phys != self.normalized_shape[0] and not carry:
            return self._forward_phys(x)
        if x.is_cuda and dispatch.ln_supported(self, x):
            from . import norm as _norm
            return _norm.layer_norm(x, self.weight, self.bias, self.eps, carry=carry)
        y = super().forward(x)
        return (y, x) if carry else y

    def forward_carry(self, x: torch.Tensor):
        """``(LN(x), x)``: feed the second value to the residual add that closes the block.  On the kernel
        path the residual's gradient is folded into the LayerNorm backward (no separate add kernel)."""
        return self.forward(x, carry=True)


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad):  # type: ignore[override]
        from ._lib import check, require_library
        from .fused import DT, ptr, stream_ptr
        lib = require_library()
        n, c, h, w = x.shape
        p, q = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
        y = torch.empty((n, p, q, c), dtype=x.dtype, device=x.device)
        idx = torch.empty((n, p, q, c), dtype=torch.uint8, device=x.device)
        check(lib.b200dp_maxpool_fwd(ptr(x), ptr(y), ptr(idx), n, h, w, c, k, stride, pad, DT[x.dtype], stream_ptr(x.device)),
              "maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.geom = (n, c, h, w, k, stride, pad)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):  # type: ignore[override]
        from ._lib import check, require_library
        from .fused import DT, ptr, stream_ptr
        lib = require_library()
        (idx,) = ctx.saved_tensors
        n, c, h, w, k, stride, pad = ctx.geom
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((n, h, w, c), dtype=dy.dtype, device=dy.device)
        check(lib.b200dp_maxpool_bwd(ptr(dy), ptr(idx), ptr(dx), n, h, w, c, k, stride, pad, DT[dy.dtype], stream_ptr(dy.device)),
              "maxpool_bwd")
        return dx.permute(0, 3, 1, 2), None, None, None


class MaxPool2d(nn.MaxPool2d):
    """Max pooling; fused NHWC kernels (1-byte argmax, gather backward) on CUDA channels-last 16-bit input."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        k, st, pd = self.kernel_size, self.stride, self.padding
        simple = all(isinstance(v, int) for v in (k, st, pd)) and self.dilation == 1 and not self.ceil_mode \
            and not self.return_indices
        if (x.is_cuda and simple and x.dim() == 4 and x.dtype in (torch.float16, torch.bfloat16) and x.shape[1] % 8 == 0
                and x.is_contiguous(memory_format=torch.channels_last) and dispatch.enabled("pool") and dispatch._lib_ok()):
            return _MaxPoolFn.apply(x, k, st, pd)
        return super().forward(x)

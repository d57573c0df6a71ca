"""Convolution front end.

1x1 / stride-1 convolutions on channels-last activations ARE matrix products
(``Y[rows, Cout] = X[rows, Cin] W[Cout, Cin]^T`` with ``rows = N*H*W``), so forward, dgrad and
wgrad run on the tcgen05 GEMM of ``csrc/gemm_sm100.cu`` without any data movement (W7 for the
1x1 case; 32 of ResNet-50's 53 convolutions).  Other shapes (3x3, 7x7, strided) go to the
implicit-GEMM kernel when it supports them and otherwise stay on cuDNN — `dispatch.conv_supported`
is the single place that decides, and DESIGN.md lists what is still on the library path.

Parity target: the convolutions inside the CNN models (Fairscale-DDP.py:74; BASELINE ResNets).
"""
from __future__ import annotations

from typing import Optional

import torch

from ._lib import library_available
from .gemm import gemm

_HALF = (torch.float16, torch.bfloat16)


def available() -> bool:
    return library_available()


def shape_supported(mod, x) -> bool:
    """1x1, stride 1, no padding, channel counts that keep TMA row pitches 16-byte aligned."""
    return (mod.kernel_size == (1, 1) and mod.stride == (1, 1) and mod.padding == (0, 0)
            and mod.in_channels % 8 == 0 and mod.out_channels % 8 == 0 and x.dim() == 4
            and x.is_contiguous(memory_format=torch.channels_last))


class _Conv1x1Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, dtype):  # type: ignore[override]
        n, cin, h, w = x.shape
        cout = weight.shape[0]
        if x.dtype != dtype:
            x = x.to(dtype)
        x2 = x.permute(0, 2, 3, 1).reshape(-1, cin)            # a view: channels-last memory is [rows, Cin]
        w16 = weight.reshape(cout, cin)
        if w16.dtype != dtype:
            w16 = w16.to(dtype)
        b32 = None if bias is None else bias.float()
        y2 = gemm(x2, w16, x2.shape[0], cout, cin, bias=b32)
        ctx.save_for_backward(x2, w16)
        ctx.dims = (n, cin, h, w, cout)
        ctx.w_shape, ctx.w_dtype, ctx.has_bias = weight.shape, weight.dtype, bias is not None
        return y2.view(n, h, w, cout).permute(0, 3, 1, 2)     # logically NCHW, physically NHWC

    @staticmethod
    def backward(ctx, dy):  # type: ignore[override]
        x2, w16 = ctx.saved_tensors
        n, cin, h, w, cout = ctx.dims
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dy2 = dy.permute(0, 2, 3, 1).reshape(-1, cout)
        if dy2.dtype != x2.dtype:
            dy2 = dy2.to(x2.dtype)
        rows = x2.shape[0]
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w16, rows, cin, cout, a_mn=False, b_mn=True).view(n, h, w, cin).permute(0, 3, 1, 2)
        if ctx.needs_input_grad[1]:
            dw = gemm(dy2, x2, cout, cin, rows, a_mn=True, b_mn=True, out_dtype=torch.float32)
            dw = dw.view(cout, 1, 1, cin).permute(0, 3, 1, 2)  # channels-last strides, like the parameter
            if ctx.w_dtype != torch.float32:
                dw = dw.to(ctx.w_dtype)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = torch.sum(dy2, dim=0, dtype=torch.float32)
        return dx, dw, db, None


def conv2d(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], stride, padding, dilation, groups):
    dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else x.dtype
    with torch.autocast("cuda", enabled=False):
        return _Conv1x1Fn.apply(x, weight, bias, dtype)

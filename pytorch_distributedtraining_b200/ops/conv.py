"""Convolution front end (W7).

On channels-last 16-bit activations:
  * 1x1 / stride-1 convolutions ARE matrix products (``Y[rows, Cout] = X[rows, Cin] W^T``): forward,
    dgrad and wgrad run on the tcgen05 GEMM (``csrc/gemm_sm100.cu``) with no data movement;
  * every other supported shape (3x3, strided 1x1, 5x5 ...; Cin % 64 == 0) runs forward — and, for
    stride 1, the input gradient — on the implicit-GEMM kernel of ``csrc/conv_sm100.cu`` (TMA fetches
    the shifted / strided pixel boxes directly from the NHWC tensor, zero-filling the padding);
  * weight gradients of all of these run on the wgrad kernel of the same file (both operands
    MN-major by 4-D TMA, pixels as the reduction, split-K with TMA reduce-add into fp32);
  * the input gradient of a stride-2 convolution is four stride-1 problems, one per parity class of the
    input pixel, each with its own tap list, run on the same dgrad kernel;
  * what is not covered stays on cuDNN and is listed in DESIGN.md: first-layer convolutions (Cin = 3).

``dispatch.conv_supported`` is the single place that decides.
Parity target: the convolutions inside the CNN models (Fairscale-DDP.py:74; BASELINE ResNets).
"""
from __future__ import annotations

from typing import Optional

import torch

from ._lib import check, library_available, require_library
from .fused import DT, ptr, stream_ptr
from .gemm import gemm

_HALF = (torch.float16, torch.bfloat16)


def available() -> bool:
    return library_available()


def _is_1x1_s1(mod) -> bool:
    return mod.kernel_size == (1, 1) and mod.stride == (1, 1) and mod.padding == (0, 0)


def smallc_supported(mod) -> bool:
    """First-layer envelope of ``csrc/conv_smallc_sm100.cu``: Cin <= 4 (RGB), stride 1 or 2, filters up to 7 x 7,
    Cout a multiple of 8 up to 64."""
    r, s = mod.kernel_size
    st = mod.stride[0]
    if not (mod.in_channels <= 4 and st in (1, 2) and mod.stride[0] == mod.stride[1] and r <= 7):
        return False
    kr = (s * (8 // st) + 15) // 16 * 16
    return (kr <= 64 and r * kr <= 512 and 8 <= mod.out_channels <= 64 and mod.out_channels % 8 == 0
            and mod.padding[0] == mod.padding[1] and isinstance(mod.padding[0], int))


def shape_supported(mod, x) -> bool:
    if x.dim() != 4:
        return False
    from . import dispatch
    if smallc_supported(mod) and dispatch.enabled("conv_first"):
        # a 3-channel NCHW image has no channels-last "contiguity" to speak of: the front end lays it out itself
        return x.shape[2] + 2 * mod.padding[0] >= mod.kernel_size[0] and x.shape[3] + 2 * mod.padding[0] >= mod.kernel_size[1]
    if not x.is_contiguous(memory_format=torch.channels_last):
        return False
    if mod.out_channels % 8 != 0:
        return False
    if _is_1x1_s1(mod):
        return mod.in_channels % 8 == 0
    return (mod.in_channels % 64 == 0 and mod.stride[0] == mod.stride[1] and mod.stride[0] in (1, 2)
            and mod.padding[0] == mod.padding[1] and isinstance(mod.padding[0], int)
            and x.shape[2] + 2 * mod.padding[0] >= mod.kernel_size[0])


def conv2d_raw(mode: int, x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], n: int, h: int, wd: int, cin: int,
               cout: int, r: int, s: int, stride: int, pad: int, colstats: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Direct call of ``b200dp_conv2d`` on NHWC-contiguous 16-bit tensors; returns NCHW-shaped,
    channels-last output.  ``colstats`` (forward only; fp32, >= 2*cout, zeroed by the caller) receives
    the per-channel sum / sum of squares of the output from the same epilogue."""
    lib = require_library()
    p, q = (h + 2 * pad - r) // stride + 1, (wd + 2 * pad - s) // stride + 1
    if mode == 0:
        out = torch.empty((n, p, q, cout), dtype=x.dtype, device=x.device)
    else:
        out = torch.empty((n, h, wd, cin), dtype=x.dtype, device=x.device)
    if colstats is not None:
        assert mode == 0 and colstats.dtype == torch.float32 and colstats.numel() >= 2 * cout
        check(lib.b200dp_conv2d_colstats(ptr(x), ptr(w), ptr(out), ptr(bias), n, h, wd, cin, cout, r, s, stride, pad,
                                         DT[x.dtype], ptr(colstats), stream_ptr(x.device)), "conv2d_colstats")
        return out.permute(0, 3, 1, 2)
    check(lib.b200dp_conv2d(mode, ptr(x), ptr(w), ptr(out), ptr(bias), n, h, wd, cin, cout, r, s, stride, pad, DT[x.dtype],
                            stream_ptr(x.device)), "conv2d")
    return out.permute(0, 3, 1, 2)


def conv2d_smallc_raw(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], n: int, h: int, wd: int, cin: int, cout: int,
                      r: int, s: int, stride: int, pad: int) -> torch.Tensor:
    """First-layer forward (Cin <= 4) on ``b200dp_conv2d_smallc_fwd``; NHWC 16-bit in, NCHW-shaped channels-last out."""
    lib = require_library()
    p, q = (h + 2 * pad - r) // stride + 1, (wd + 2 * pad - s) // stride + 1
    out = torch.empty((n, p, q, cout), dtype=x.dtype, device=x.device)
    check(lib.b200dp_conv2d_smallc_fwd(ptr(x), ptr(w), ptr(out), ptr(bias), n, h, wd, cin, cout, r, s, stride, pad, DT[x.dtype],
                                       stream_ptr(x.device)), "conv2d_smallc_fwd")
    return out.permute(0, 3, 1, 2)


def conv2d_smallc_wgrad_raw(x: torch.Tensor, dy: torch.Tensor, n: int, h: int, wd: int, cin: int, cout: int, r: int, s: int,
                            stride: int, pad: int) -> torch.Tensor:
    lib = require_library()
    dw = torch.zeros((cout, r, s, cin), dtype=torch.float32, device=x.device)       # the kernel accumulates atomically
    check(lib.b200dp_conv2d_smallc_wgrad(ptr(x), ptr(dy), ptr(dw), n, h, wd, cin, cout, r, s, stride, pad, DT[x.dtype],
                                         stream_ptr(x.device)), "conv2d_smallc_wgrad")
    return dw.permute(0, 3, 1, 2)


def conv2d_wgrad_raw(x: torch.Tensor, dy: torch.Tensor, n: int, h: int, wd: int, cin: int, cout: int, r: int, s: int,
                     stride: int, pad: int) -> torch.Tensor:
    """fp32 weight gradient in the channels-last weight layout ([Cout][R][S][Cin] memory)."""
    lib = require_library()
    dw = torch.empty((cout, r, s, cin), dtype=torch.float32, device=x.device)
    check(lib.b200dp_conv2d_wgrad(ptr(x), ptr(dy), ptr(dw), n, h, wd, cin, cout, r, s, stride, pad, DT[x.dtype], 0,
                                  stream_ptr(x.device)), "conv2d_wgrad")
    return dw.permute(0, 3, 1, 2)


class _ConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, dtype, stride, pad, colstats=None, carry=False):  # type: ignore[override]
        x_in = x
        n, cin, h, w = x.shape
        cout, _, r, s = weight.shape
        if x.dtype != dtype:
            x = x.to(dtype)
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        from .fused import lp_weight
        w16 = lp_weight(weight, dtype)        # the optimizer's 16-bit shadow when it has one: no cast kernel
        if not w16.is_contiguous(memory_format=torch.channels_last):
            w16 = w16.contiguous(memory_format=torch.channels_last)       # memory = [Cout][R][S][Cin]
        b32 = None if bias is None else bias.float()
        one = r == 1 and s == 1 and stride == 1 and pad == 0
        small = cin <= 4
        if small:
            colstats = None
            y = conv2d_smallc_raw(x, w16, b32, n, h, w, cin, cout, r, s, stride, pad)
        elif one:
            x2 = x.permute(0, 2, 3, 1).reshape(-1, cin)
            if colstats is not None and b32 is None:
                y = gemm(x2, w16.reshape(cout, cin), x2.shape[0], cout, cin, colstats=colstats)
            else:
                y = gemm(x2, w16.reshape(cout, cin), x2.shape[0], cout, cin, bias=b32)
                colstats = None
            y = y.view(n, h, w, cout).permute(0, 3, 1, 2)
        else:
            y = conv2d_raw(0, x, w16, b32, n, h, w, cin, cout, r, s, stride, pad, colstats=colstats)
        ctx.emitted_stats = colstats is not None
        ctx.save_for_backward(x, w16)
        ctx.geom = (n, cin, h, w, cout, r, s, stride, pad, one)
        ctx.small = small
        ctx.w_dtype, ctx.has_bias = weight.dtype, bias is not None
        if carry:
            # second output: the input itself, for the branch that bypasses this convolution (residual
            # identity / downsample).  Its gradient comes back into THIS backward, where the input
            # gradient is accumulated straight into it — autograd never runs a separate add kernel.
            return y, x_in.view_as(x_in)
        return y

    @staticmethod
    def backward(ctx, dy, d_carry=None):  # type: ignore[override]
        x, w16 = ctx.saved_tensors
        n, cin, h, w, cout, r, s, stride, pad, one = ctx.geom
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = dw = db = None
        need_dx, need_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        if ctx.small:
            if need_dw:
                dw = conv2d_smallc_wgrad_raw(x, dy, n, h, w, cin, cout, r, s, stride, pad)
            if need_dx:
                from . import dispatch
                if cout % 64 == 0 and (stride == 1 or (stride == 2 and dispatch.enabled("dgrad_s2"))):
                    # rare (perceptual loss on an RGB output): pad Cin to 8 and use the regular input-gradient kernel
                    w8 = torch.zeros((cout, 8, r, s), dtype=w16.dtype, device=w16.device).contiguous(memory_format=torch.channels_last)
                    w8[:, :cin] = w16
                    dx = conv2d_raw(1, dy, w8, None, n, h, w, 8, cout, r, s, stride, pad)[:, :cin]
                else:
                    dx = torch.ops.aten.convolution_backward(dy, x, w16, None, (stride, stride), (pad, pad), (1, 1), False,
                                                             (0, 0), 1, (True, False, False))[0]
        elif one:
            x2 = x.permute(0, 2, 3, 1).reshape(-1, cin)
            dy2 = dy.permute(0, 2, 3, 1).reshape(-1, cout)
            rows = x2.shape[0]
            # weight gradient first: the input gradient is then the most recent tensor in L2 when the
            # backward of the layer below (BatchNorm) starts reading it
            if need_dw:
                dw = gemm(dy2, x2, cout, cin, rows, a_mn=True, b_mn=True, out_dtype=torch.float32)
                dw = dw.view(cout, 1, 1, cin).permute(0, 3, 1, 2)
            if need_dx:
                acc = _accumulable(d_carry, x)
                if acc:
                    # dX += dY W, added onto the bypass branch's gradient by the epilogue (TMA reduce-add)
                    gemm(dy2, w16.reshape(cout, cin), rows, cin, cout, a_mn=False, b_mn=True,
                         out=d_carry.permute(0, 2, 3, 1).reshape(rows, cin), accumulate=True)
                    dx, d_carry = d_carry, None
                else:
                    dx = gemm(dy2, w16.reshape(cout, cin), rows, cin, cout, a_mn=False, b_mn=True).view(n, h, w, cin).permute(0, 3, 1, 2)
        else:
            from . import dispatch
            ours_dx = need_dx and (stride == 1 or (stride == 2 and dispatch.enabled("dgrad_s2"))) and cin % 8 == 0
            if need_dw:
                dw = conv2d_wgrad_raw(x, dy, n, h, w, cin, cout, r, s, stride, pad)
            if ours_dx and cout % 64 != 0:
                # the input-gradient kernel reduces over Cout in 64-channel chunks: zero-pad dY and the filters to the next
                # multiple of 64 (zeros add nothing) instead of handing the layer to the library
                cp = (cout + 63) // 64 * 64
                dyp = torch.cat([dy, dy.new_zeros((n, cp - cout, dy.shape[2], dy.shape[3]))], dim=1).contiguous(memory_format=torch.channels_last)
                wp = torch.cat([w16, w16.new_zeros((cp - cout, cin, r, s))], dim=0).contiguous(memory_format=torch.channels_last)
                dx = conv2d_raw(1, dyp, wp, None, n, h, w, cin, cp, r, s, stride, pad)
            elif ours_dx:
                dx = conv2d_raw(1, dy, w16, None, n, h, w, cin, cout, r, s, stride, pad)
            if need_dx and not ours_dx:
                # library path for what the implicit-GEMM kernels do not cover: input gradients
                # whose reduction width (Cout) is not a multiple of 64
                dx = torch.ops.aten.convolution_backward(dy, x, w16, None, (stride, stride), (pad, pad), (1, 1), False,
                                                         (0, 0), 1, (True, False, False))[0]
        if d_carry is not None:
            dx = d_carry if dx is None else dx + d_carry.to(dx.dtype)
        if dw is not None and dw.dtype != ctx.w_dtype:
            dw = dw.to(ctx.w_dtype)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = torch.sum(dy, dim=(0, 2, 3), dtype=torch.float32)
        return dx, dw, db, None, None, None, None, None


def _accumulable(d_carry, x) -> bool:
    """The bypass gradient can take the input gradient in place: same dtype / shape, dense NHWC, not a view."""
    return (d_carry is not None and d_carry.dtype == x.dtype and d_carry.shape == x.shape and d_carry._base is None
            and d_carry.is_contiguous(memory_format=torch.channels_last) and d_carry.data_ptr() % 16 == 0)


BN_STATS_ATTR = "_b200dp_bn_sums"


class _StatsArena:
    """Zeroed fp32 scratch for the conv -> BatchNorm statistics, handed out in slices.

    Every statistics vector has to start at zero (the epilogues accumulate into it), which used to be one fill launch
    per convolution and step — 53 for ResNet-50.  Here the vectors of a whole optimizer step are slices of one block that
    is zeroed ONCE, at the first request after the parameters changed (``ops.fused.PARAM_EPOCH``: bumped by the fused
    optimizers, the ZeRO engine and every graph replay that contains an optimizer step).  Without such a signal (a
    plain torch optimizer, forward-only loops) the block is simply used up and a fresh zeroed one is allocated: one
    memset per ~quarter million floats instead of one per layer.  Under CUDA-graph capture the same logic is recorded:
    the graph of the first micro-batch contains the block's memset, later micro-batches of the step take the slices
    behind it."""

    BLOCK = 1 << 18          # floats

    def __init__(self, device: torch.device):
        self.device, self.buf, self.cursor, self.epoch = device, None, 0, -1

    def take(self, n: int) -> torch.Tensor:
        from . import fused
        n_al = (n + 63) // 64 * 64
        if self.buf is None or self.cursor + n_al > self.buf.numel():
            self.buf = torch.zeros(max(self.BLOCK, n_al), dtype=torch.float32, device=self.device)
            self.cursor, self.epoch = 0, fused.PARAM_EPOCH
        elif self.epoch != fused.PARAM_EPOCH:
            self.buf.zero_()
            self.cursor, self.epoch = 0, fused.PARAM_EPOCH
        out = self.buf[self.cursor: self.cursor + n]
        self.cursor += n_al
        return out


_STATS_ARENAS: dict = {}


def _stats_vector(n: int, device: torch.device) -> torch.Tensor:
    arena = _STATS_ARENAS.get(device)
    if arena is None:
        arena = _STATS_ARENAS[device] = _StatsArena(device)
    return arena.take(n)


def conv2d(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], stride, padding, dilation, groups,
           emit_bn_stats: bool = False, carry: bool = False):
    """``emit_bn_stats``: the convolution's epilogue also produces [sum(y), sum(y^2), (count slot)] per
    output channel and hangs that vector on the returned tensor; the BatchNorm that consumes this very
    tensor (``ops.norm.batch_norm_act``) then skips its own reduction pass over the activation.
    ``carry``: also return the input (for a bypass branch); see ``_ConvFn.forward``."""
    dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else x.dtype
    with torch.autocast("cuda", enabled=False):
        sums = None
        if emit_bn_stats and (bias is None or tuple(weight.shape[2:]) != (1, 1) or int(stride[0]) != 1):
            sums = _stats_vector(2 * weight.shape[0] + 1, x.device)
        out = _ConvFn.apply(x, weight, bias, dtype, int(stride[0]), int(padding[0]), sums, carry)
        y = out[0] if carry else out
        if sums is not None:
            setattr(y, BN_STATS_ATTR, sums)
        return out

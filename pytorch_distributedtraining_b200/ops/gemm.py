"""tcgen05 / TMEM / TMA GEMM front end (``csrc/gemm_sm100.cu``): raw ``gemm`` and the
``linear`` autograd function built on it (forward, dgrad, wgrad on the same kernel with
different operand majors — no transposes, no cuBLAS).

Parity target: the ``nn.Linear`` layers inside SwinIR / ViT (Stoke-DDP.py:206-208) and, through
``ops.conv``, the 1x1 convolutions of the CNN models.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.nn.functional as F

from ._lib import check, library_available, require_library
from .fused import DT, ptr, stream_ptr

_HALF = (torch.float16, torch.bfloat16)


def available() -> bool:
    return library_available()


def gemm(a: torch.Tensor, b: torch.Tensor, M: int, N: int, K: int, a_mn: bool = False, b_mn: bool = False,
         bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None, out_dtype: Optional[torch.dtype] = None,
         accumulate: bool = False, colstats: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``C[M,N] (+)= A(MxK) @ B(NxK)^T (+ bias)``.  ``a`` / ``b`` are 2-D, row-major, 16-bit:
    K-major operands are ``[rows, K]``, MN-major operands are ``[K, rows]``.  ``colstats`` (fp32,
    >= 2N, zeroed by the caller): the epilogue also adds the column sums / sums of squares of C into
    it (the batch statistics for a BatchNorm that follows a 1x1 convolution)."""
    lib = require_library()
    assert a.dtype in _HALF and b.dtype == a.dtype and a.dim() == 2 and b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1
    out_dtype = out_dtype or a.dtype
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    assert out.stride(1) == 1 and out.shape == (M, N)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.is_contiguous()
    if colstats is not None:
        assert not (a_mn or b_mn or accumulate) and bias is None and out.dtype == a.dtype
        assert colstats.dtype == torch.float32 and colstats.numel() >= 2 * N and colstats.is_contiguous()
        check(lib.b200dp_gemm_colstats(ptr(a), ptr(b), ptr(out), M, N, K, a.stride(0), b.stride(0), out.stride(0),
                                       DT[a.dtype], ptr(colstats), stream_ptr(a.device)), "gemm_colstats")
        return out
    check(lib.b200dp_gemm(ptr(a), ptr(b), ptr(out), ptr(bias), M, N, K, a.stride(0), b.stride(0), out.stride(0),
                          int(a_mn), int(b_mn), DT[a.dtype], DT[out.dtype], int(accumulate), stream_ptr(a.device)),
          "gemm")
    return out


class _LinearFn(torch.autograd.Function):
    """y = x W^T + b with 16-bit compute; ``weight`` / ``bias`` are the fp32 parameters, so dW
    and db come out in fp32 directly (no cast kernels on the gradient path)."""

    @staticmethod
    def forward(ctx, x, weight, bias, dtype):  # type: ignore[override]
        in_f = weight.shape[1]
        x2 = x.reshape(-1, in_f)
        if x2.dtype != dtype:
            x2 = x2.to(dtype)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        w16 = weight if weight.dtype == dtype else weight.to(dtype)
        if not w16.is_contiguous():
            w16 = w16.contiguous()
        b32 = None if bias is None else (bias if bias.dtype == torch.float32 else bias.float())
        rows, out_f = x2.shape[0], weight.shape[0]
        y = gemm(x2, w16, rows, out_f, in_f, bias=b32)
        ctx.save_for_backward(x2, w16)
        ctx.x_shape, ctx.has_bias, ctx.w_dtype = x.shape, bias is not None, weight.dtype
        ctx.b_dtype = None if bias is None else bias.dtype
        return y.view(*x.shape[:-1], out_f)

    @staticmethod
    def backward(ctx, dy):  # type: ignore[override]
        x2, w16 = ctx.saved_tensors
        rows, in_f = x2.shape
        out_f = w16.shape[0]
        dy2 = dy.reshape(rows, out_f)
        if dy2.dtype != x2.dtype:
            dy2 = dy2.to(x2.dtype)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = None
        # parameter gradients first, the input gradient last: it is then the most recent tensor in L2 for
        # the backward of the layer below
        if ctx.needs_input_grad[1]:
            dw = gemm(dy2, x2, out_f, in_f, rows, a_mn=True, b_mn=True, out_dtype=torch.float32)
            if ctx.w_dtype != torch.float32:
                dw = dw.to(ctx.w_dtype)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            if out_f % 8 == 0 and out_f <= 16384 and rows >= 64:
                from .norm import column_sums
                db = column_sums(dy2)
            else:
                db = torch.sum(dy2, dim=0, dtype=torch.float32)
            if ctx.b_dtype != torch.float32:
                db = db.to(ctx.b_dtype)
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w16, rows, in_f, out_f, a_mn=False, b_mn=True).view(ctx.x_shape)
        return dx, dw, db, None


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], act: Optional[str] = None) -> torch.Tensor:
    dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else x.dtype
    if dtype not in _HALF:
        y = F.linear(x, weight, bias)
    else:
        # the function does its own casting; keep autocast from re-casting the fp32 parameters
        with torch.autocast("cuda", enabled=False):
            y = _LinearFn.apply(x, weight, bias, dtype)
    if act == "gelu":
        y = F.gelu(y)
    elif act == "relu":
        y = F.relu(y)
    return y


_GATHER_FLAGS: dict = {}


def gemm_gather(x: torch.Tensor, shard_ptrs, w_full: torch.Tensor, rank: int, world: int, rows_per_rank: int,
                bias: Optional[torch.Tensor] = None, ctx=None, tile_flags: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``Y = X @ W.T (+ bias)`` with ``W`` row-sharded over ranks: the all-gather of ``W`` happens inside
    the GEMM kernel (W4).  ``shard_ptrs``: per-rank shard addresses as mapped here (ints or tensors);
    ``w_full``: local ``[N, K]`` buffer the kernel fills; ``ctx``: ``ops.collectives.PeerContext`` or
    ``None`` when every shard is local."""
    from .collectives import table
    lib = require_library()
    M, K = x.shape
    N = w_full.shape[0]
    assert x.dtype in _HALF and w_full.dtype == x.dtype and x.stride(1) == 1 and w_full.is_contiguous()
    y = torch.empty((M, N), dtype=x.dtype, device=x.device)
    if tile_flags is None:
        # arrival counters, one set per (device, stream): zero at creation, the kernel leaves them zero
        key = (x.device.index, torch.cuda.current_stream(x.device).cuda_stream, (N + 127) // 128 + 1)
        tile_flags = _GATHER_FLAGS.get(key)
        if tile_flags is None:
            tile_flags = _GATHER_FLAGS[key] = torch.zeros(key[2], dtype=torch.int32, device=x.device)
    check(lib.b200dp_gemm_gather(ptr(x), table(shard_ptrs), ptr(w_full), ptr(y), ptr(bias), M, N, K, x.stride(0), y.stride(0),
                                 world, rank, rows_per_rank, DT[x.dtype], ptr(tile_flags),
                                 ctypes.byref(ctx.c) if ctx is not None else None, stream_ptr(x.device)), "gemm_gather")
    return y

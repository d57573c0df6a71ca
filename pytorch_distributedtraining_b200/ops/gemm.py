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


def fused_supported(M: int, N: int) -> bool:
    """Is ``[M, N]`` inside the envelope of the CTA-pair kernel's fused epilogues (``b200dp_gemm_fused``)?"""
    return bool(require_library().b200dp_gemm_fused_supported(int(M), int(N)))


def gemm_fused(a: torch.Tensor, b: torch.Tensor, M: int, N: int, K: int, mode: int, bias: Optional[torch.Tensor] = None,
               aux: Optional[torch.Tensor] = None):
    """Large-product GEMM with a fused epilogue (``include/b200dp.h``): mode 1 returns ``(C, gelu(C))``,
    mode 2 ``(A B) * gelu'(aux)`` with ``b`` given MN-major (``[K, N]``, the dgrad form), mode 3
    ``A B^T (+bias) + aux``."""
    lib = require_library()
    assert a.dtype in _HALF and b.dtype == a.dtype and a.stride(1) == 1 and b.stride(1) == 1
    c = torch.empty((M, N), dtype=a.dtype, device=a.device)
    c2 = torch.empty((M, N), dtype=a.dtype, device=a.device) if mode == 1 else None
    if aux is not None:
        assert aux.dtype == a.dtype and aux.shape == (M, N) and aux.stride(1) == 1
    check(lib.b200dp_gemm_fused(ptr(a), ptr(b), ptr(c), ptr(c2), ptr(bias), ptr(aux), M, N, K, a.stride(0), b.stride(0),
                                c.stride(0), N, aux.stride(0) if aux is not None else 0, int(mode == 2), DT[a.dtype], mode,
                                stream_ptr(a.device)), "gemm_fused")
    return (c, c2) if mode == 1 else c


def _as_rows(t: torch.Tensor, cols: int, dtype: torch.dtype) -> torch.Tensor:
    t2 = t.reshape(-1, cols)
    if t2.dtype != dtype:
        t2 = t2.to(dtype)
    return t2 if t2.is_contiguous() else t2.contiguous()


def _w16(w: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    from .fused import lp_weight
    w = lp_weight(w, dtype)                 # the optimizer's 16-bit shadow when it has one: no cast kernel
    return w if w.is_contiguous() else w.contiguous()


def _bias_grad(dy2: torch.Tensor) -> torch.Tensor:
    rows, out_f = dy2.shape
    if out_f % 8 == 0 and out_f <= 16384 and rows >= 64:
        from .norm import column_sums
        return column_sums(dy2)
    return torch.sum(dy2, dim=0, dtype=torch.float32)


class _LinearFn(torch.autograd.Function):
    """y = x W^T + b (+ residual) with 16-bit compute; ``weight`` / ``bias`` are the fp32 parameters, so dW
    and db come out in fp32 directly (no cast kernels on the gradient path).  The residual add is part of
    the GEMM epilogue when the shape is inside the CTA-pair kernel's envelope."""

    @staticmethod
    def forward(ctx, x, weight, bias, dtype, residual=None):  # type: ignore[override]
        in_f = weight.shape[1]
        x2 = _as_rows(x, in_f, dtype)
        w16 = _w16(weight, dtype)
        b32 = None if bias is None else (bias if bias.dtype == torch.float32 else bias.float())
        rows, out_f = x2.shape[0], weight.shape[0]
        if residual is not None and fused_supported(rows, out_f):
            y = gemm_fused(x2, w16, rows, out_f, in_f, 3, bias=b32, aux=_as_rows(residual, out_f, dtype))
        else:
            y = gemm(x2, w16, rows, out_f, in_f, bias=b32)
            if residual is not None:
                y += residual.reshape(rows, out_f)
        ctx.save_for_backward(x2, w16)
        ctx.x_shape, ctx.has_bias, ctx.w_dtype = x.shape, bias is not None, weight.dtype
        ctx.b_dtype = None if bias is None else bias.dtype
        ctx.res_meta = None if residual is None else (residual.shape, residual.dtype)
        return y.view(*x.shape[:-1], out_f)

    @staticmethod
    def backward(ctx, dy):  # type: ignore[override]
        x2, w16 = ctx.saved_tensors
        rows, in_f = x2.shape
        out_f = w16.shape[0]
        dy2 = _as_rows(dy, out_f, x2.dtype)
        dx = dw = db = dres = None
        # parameter gradients first, the input gradient last: it is then the most recent tensor in L2 for
        # the backward of the layer below
        if ctx.needs_input_grad[1]:
            dw = gemm(dy2, x2, out_f, in_f, rows, a_mn=True, b_mn=True, out_dtype=torch.float32)
            if ctx.w_dtype != torch.float32:
                dw = dw.to(ctx.w_dtype)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = _bias_grad(dy2)
            if ctx.b_dtype != torch.float32:
                db = db.to(ctx.b_dtype)
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w16, rows, in_f, out_f, a_mn=False, b_mn=True).view(ctx.x_shape)
        if ctx.res_meta is not None and ctx.needs_input_grad[4]:
            dres = dy.reshape(ctx.res_meta[0])          # the residual's gradient IS dy (no copy)
            if dres.dtype != ctx.res_meta[1]:
                dres = dres.to(ctx.res_meta[1])
        return dx, dw, db, None, dres


class _MLPFn(torch.autograd.Function):
    """``y = fc2(gelu(fc1(x))) (+ residual)`` of a transformer block as ONE autograd node over four GEMM
    launches forward and five backward, with every elementwise step inside a GEMM epilogue:
    fc1 writes the pre-activation and its GELU from the same epilogue, fc2 adds the residual in its
    epilogue, and in the backward the input gradient of fc2 is multiplied by gelu'(pre-activation) on
    its way out of the tensor core.  (Reference: the MLP of the ViT / SwinIR blocks, Stoke-DDP.py:206-208;
    Linear, GELU and add are separate kernels there.)"""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, dtype, residual=None):  # type: ignore[override]
        in_f, hid, out_f = w1.shape[1], w1.shape[0], w2.shape[0]
        x2 = _as_rows(x, in_f, dtype)
        rows = x2.shape[0]
        w1h, w2h = _w16(w1, dtype), _w16(w2, dtype)
        b1f = None if b1 is None else b1.float()
        b2f = None if b2 is None else b2.float()
        z, a = gemm_fused(x2, w1h, rows, hid, in_f, 1, bias=b1f)
        if residual is not None and fused_supported(rows, out_f):
            y = gemm_fused(a, w2h, rows, out_f, hid, 3, bias=b2f, aux=_as_rows(residual, out_f, dtype))
        else:
            y = gemm(a, w2h, rows, out_f, hid, bias=b2f)
            if residual is not None:
                y += residual.reshape(rows, out_f)
        ctx.save_for_backward(x2, z, a, w1h, w2h)
        ctx.meta = (x.shape, w1.dtype, w2.dtype, None if b1 is None else b1.dtype, None if b2 is None else b2.dtype,
                    None if residual is None else (residual.shape, residual.dtype))
        return y.view(*x.shape[:-1], out_f)

    @staticmethod
    def backward(ctx, dy):  # type: ignore[override]
        x2, z, a, w1h, w2h = ctx.saved_tensors
        x_shape, w1_dt, w2_dt, b1_dt, b2_dt, res_meta = ctx.meta
        rows, in_f = x2.shape
        hid, out_f = w1h.shape[0], w2h.shape[0]
        dy2 = _as_rows(dy, out_f, x2.dtype)
        dw2 = gemm(dy2, a, out_f, hid, rows, a_mn=True, b_mn=True, out_dtype=torch.float32)
        db2 = None if b2_dt is None else _bias_grad(dy2).to(b2_dt)
        # dz = (dy W2) * gelu'(z), straight out of the dgrad epilogue
        dz = gemm_fused(dy2, w2h, rows, hid, out_f, 2, aux=z)
        dw1 = gemm(dz, x2, hid, in_f, rows, a_mn=True, b_mn=True, out_dtype=torch.float32)
        db1 = None if b1_dt is None else _bias_grad(dz).to(b1_dt)
        dx = gemm(dz, w1h, rows, in_f, hid, a_mn=False, b_mn=True).view(x_shape) if ctx.needs_input_grad[0] else None
        dres = None
        if res_meta is not None and ctx.needs_input_grad[6]:
            dres = dy.reshape(res_meta[0])
            if dres.dtype != res_meta[1]:
                dres = dres.to(res_meta[1])
        return dx, dw1.to(w1_dt), db1, dw2.to(w2_dt), db2, None, dres


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], act: Optional[str] = None,
           residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else x.dtype
    if dtype not in _HALF:
        y = F.linear(x, weight, bias)
    else:
        # the function does its own casting; keep autocast from re-casting the fp32 parameters
        with torch.autocast("cuda", enabled=False):
            y = _LinearFn.apply(x, weight, bias, dtype, residual if act is None else None)
        if act is None:
            return y
    if act == "gelu":
        y = F.gelu(y)
    elif act == "relu":
        y = F.relu(y)
    return y if residual is None else y + residual


def mlp_supported(x: torch.Tensor, w1: torch.Tensor, w2: torch.Tensor) -> bool:
    """Fused transformer MLP: 16-bit compute on CUDA and both products inside the CTA-pair envelope."""
    from . import dispatch
    if not (x.is_cuda and dispatch.enabled("gemm") and dispatch.enabled("mlp") and dispatch._lib_ok()):
        return False
    dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else x.dtype
    rows = x.numel() // x.shape[-1]
    return (dtype in _HALF and all(d % 8 == 0 for d in (*w1.shape, *w2.shape))
            and fused_supported(rows, w1.shape[0]) and fused_supported(rows, w1.shape[1]))


def mlp(x: torch.Tensor, w1: torch.Tensor, b1: Optional[torch.Tensor], w2: torch.Tensor, b2: Optional[torch.Tensor],
        residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``fc2(gelu(fc1(x))) (+ residual)``; caller checks ``mlp_supported`` first."""
    dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else x.dtype
    with torch.autocast("cuda", enabled=False):
        return _MLPFn.apply(x, w1, b1, w2, b2, dtype, residual)


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

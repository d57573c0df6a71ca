"""Layout glue around the attention core of the transformer models (ViT-B/16, SwinIR).

The attention product itself is the library flash kernel (``F.scaled_dot_product_attention``; DESIGN.md
lists it under "still on a library kernel").  What this module owns is everything between the packed
``[B, S, 3*D]`` projection and that kernel, forward and backward:

  * ``split_qkv``: q / k / v as strided per-head views of the projection output (no copy); the backward
    packs the three per-head gradients into the token-major ``[B, S, 3, H, d]`` gradient of the
    projection with ONE streaming kernel (``b200dp_heads_to_tokens``) instead of stack + strided copy;
  * ``merge_heads``: ``[B, H, S, d] -> [B, S, D]``: a view when the attention kernel already wrote
    token-major memory, one streaming kernel otherwise; the backward is a view.

Parity target: the reshape / permute chain inside ``nn.MultiheadAttention`` as used by the reference's
transformer model (Stoke-DDP.py:206-208).
"""
from __future__ import annotations

import ctypes
from typing import Tuple

import torch

from . import dispatch
from ._lib import check, require_library
from .collectives import table
from .fused import ptr, stream_ptr

_HALF = (torch.float16, torch.bfloat16)


def _kernels_ok(t: torch.Tensor, d: int) -> bool:
    return t.is_cuda and t.dtype in _HALF and (d * t.element_size()) % 16 == 0 and dispatch.enabled("attn_layout") \
        and dispatch._lib_ok()


def _packable(srcs, d: int) -> bool:
    """Same strides on every source, contiguous last dimension, 16-byte aligned rows."""
    s0 = srcs[0]
    epv = 16 // s0.element_size()
    return all(t.stride() == s0.stride() and t.stride(3) == 1 and t.data_ptr() % 16 == 0 for t in srcs) \
        and all(st % epv == 0 for st in s0.stride()[:3])


def _heads_to_tokens(srcs, B: int, H: int, S: int, d: int) -> torch.Tensor:
    """W ``[B, H, S, d]`` tensors (any strides with a contiguous last dimension) -> ``[B, S, W, H, d]``."""
    lib = require_library()
    W = len(srcs)
    out = torch.empty((B, S, W, H, d), dtype=srcs[0].dtype, device=srcs[0].device)
    sb, sh, ss, _ = srcs[0].stride()
    check(lib.b200dp_heads_to_tokens(table(list(srcs)), ptr(out), B, H, S, d, W, srcs[0].element_size(),
                                     ctypes.c_longlong(sb), ctypes.c_longlong(sh), ctypes.c_longlong(ss),
                                     stream_ptr(out.device)), "heads_to_tokens")
    return out


def _tokens_to_heads(x: torch.Tensor, B: int, H: int, S: int, d: int, W: int):
    """``[B, S, W, H, d]`` (contiguous) -> W contiguous ``[B, H, S, d]`` tensors."""
    lib = require_library()
    outs = [torch.empty((B, H, S, d), dtype=x.dtype, device=x.device) for _ in range(W)]
    check(lib.b200dp_tokens_to_heads(ptr(x), table(outs), B, H, S, d, W, x.element_size(), stream_ptr(x.device)),
          "tokens_to_heads")
    return outs


class _SplitQKV(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv: torch.Tensor, heads: int):  # type: ignore[override]
        B, S, D3 = qkv.shape
        d = D3 // (3 * heads)
        ctx.geom = (B, S, heads, d)
        q, k, v = qkv.view(B, S, 3, heads, d).permute(2, 0, 3, 1, 4).unbind(0)
        return q, k, v

    @staticmethod
    def backward(ctx, dq, dk, dv):  # type: ignore[override]
        B, S, H, d = ctx.geom
        grads = (dq, dk, dv)
        if _kernels_ok(dq, d) and _packable(grads, d):
            return _heads_to_tokens(grads, B, H, S, d).view(B, S, 3 * H * d), None
        # token-major or arbitrary gradients: one strided copy per tensor into the packed buffer
        out = torch.empty((B, S, 3, H, d), dtype=dq.dtype, device=dq.device)
        for i, g in enumerate(grads):
            out[:, :, i].copy_(g.permute(0, 2, 1, 3))
        return out.view(B, S, 3 * H * d), None


class _MergeHeads(torch.autograd.Function):
    @staticmethod
    def forward(ctx, o: torch.Tensor):  # type: ignore[override]
        B, H, S, d = o.shape
        ctx.geom = (B, H, S, d)
        if o.transpose(1, 2).is_contiguous():
            ctx.mode = "view"
            return o.transpose(1, 2).reshape(B, S, H * d)
        if _kernels_ok(o, d) and _packable((o,), d):
            ctx.mode = "kernel"
            return _heads_to_tokens((o,), B, H, S, d).view(B, S, H * d)
        ctx.mode = "copy"
        return o.transpose(1, 2).reshape(B, S, H * d)

    @staticmethod
    def backward(ctx, dy):  # type: ignore[override]
        B, H, S, d = ctx.geom
        if ctx.mode == "kernel" and dy.is_contiguous():
            return _tokens_to_heads(dy, B, H, S, d, 1)[0]
        return dy.reshape(B, S, H, d).transpose(1, 2)


def split_qkv(qkv: torch.Tensor, heads: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """``[B, S, 3*D]`` projection -> q, k, v ``[B, H, S, d]`` (strided views)."""
    if not qkv.is_contiguous():
        qkv = qkv.contiguous()
    return _SplitQKV.apply(qkv, heads)


def merge_heads(o: torch.Tensor) -> torch.Tensor:
    """``[B, H, S, d]`` attention output -> ``[B, S, H*d]``."""
    return _MergeHeads.apply(o)

"""Attention for the transformer models (ViT-B/16, SwinIR).

``attention(qkv, heads, ...)`` is the tcgen05 kernel pair of ``csrc/attn_sm100.cu``: it reads q / k / v straight
out of the packed ``[B, S, 3*D]`` projection, writes the token-major ``[B, S, D]`` tensor the out projection
consumes, and its backward writes the packed ``[B, S, 3*D]`` gradient — no per-head copies in either direction.
Shapes outside its envelope (more than 256 tokens, head dimension above 64) use the library kernel
(``F.scaled_dot_product_attention``) with the layout helpers below:

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
from typing import Optional, Tuple

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


def attention_supported(qkv: torch.Tensor, heads: int, bias: Optional[torch.Tensor] = None) -> bool:
    """Envelope of the tcgen05 attention kernels: CUDA fp16 / bf16, packed ``[B, S, 3 * heads * d]`` projection,
    at most 256 tokens, even head dimension <= 64; a learnable additive bias needs S <= 64 (window attention)."""
    if not (qkv.is_cuda and qkv.dtype in _HALF and qkv.dim() == 3 and dispatch.enabled("attn") and dispatch._lib_ok()):
        return False
    B, S, D3 = qkv.shape
    if D3 % (3 * heads) != 0:
        return False
    d = D3 // (3 * heads)
    if not (0 < S <= 256 and 2 <= d <= 64 and d % 2 == 0):
        return False
    if bias is not None and bias.requires_grad and (S > 64 or dispatch.deterministic()):
        return False            # the bias gradient is accumulated with float atomics
    return True


class _AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, heads, scale, bias, mask):  # type: ignore[override]
        from .fused import DT
        lib = require_library()
        if not qkv.is_contiguous():
            qkv = qkv.contiguous()
        B, S, D3 = qkv.shape
        d = D3 // (3 * heads)
        out = torch.empty((B, S, heads * d), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((B, heads, S), dtype=torch.float32, device=qkv.device)
        b32 = None if bias is None else bias.detach().float().contiguous()        # [H, S, S]
        m32 = None if mask is None else mask.detach().float().contiguous()        # [nW, S, S]
        nW = 0 if m32 is None else m32.shape[0]
        check(lib.b200dp_attention_fwd(ptr(qkv), ptr(out), ptr(lse), ptr(b32), ptr(m32), nW, B, heads, S, d,
                                       ctypes.c_float(scale), DT[qkv.dtype], stream_ptr(qkv.device)), "attention_fwd")
        ctx.save_for_backward(qkv, out, lse, b32, m32)
        ctx.geom = (B, heads, S, d, scale, nW)
        ctx.bias_dtype = None if bias is None else bias.dtype
        return out

    @staticmethod
    def backward(ctx, dout):  # type: ignore[override]
        from .fused import DT
        lib = require_library()
        qkv, out, lse, b32, m32 = ctx.saved_tensors
        B, H, S, d, scale, nW = ctx.geom
        if dout.dtype != qkv.dtype:
            dout = dout.to(qkv.dtype)
        if not dout.is_contiguous():
            dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        dbias = None
        if b32 is not None and ctx.needs_input_grad[3]:
            dbias = torch.zeros((H, S, S), dtype=torch.float32, device=qkv.device)
        check(lib.b200dp_attention_bwd(ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(dqkv), ptr(b32), ptr(m32), nW,
                                       ptr(dbias), B, H, S, d, ctypes.c_float(scale), DT[qkv.dtype],
                                       stream_ptr(qkv.device)), "attention_bwd")
        if dbias is not None and dbias.dtype != ctx.bias_dtype:
            dbias = dbias.to(ctx.bias_dtype)
        return dqkv, None, None, dbias, None


def attention(qkv: torch.Tensor, heads: int, scale: Optional[float] = None, bias: Optional[torch.Tensor] = None,
              mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``softmax(q k^T * scale + bias + mask) v`` over the packed projection ``qkv [B, S, 3 * heads * d]``
    (memory order ``[B][S][3][heads][d]``); returns ``[B, S, heads * d]``.

    ``bias``: ``[heads, S, S]`` (may require grad: the relative-position bias of window attention);
    ``mask``: ``[nW, S, S]`` constant, applied to batch entry ``b`` as ``mask[b % nW]`` (shifted windows).
    Call ``attention_supported`` first; outside the envelope use ``attention_reference``."""
    d = qkv.shape[-1] // (3 * heads)
    return _AttentionFn.apply(qkv, heads, float(scale if scale is not None else d ** -0.5), bias, mask)


def window_attention_supported(qkv: torch.Tensor, heads: int, d: int, ws: int, bias: Optional[torch.Tensor] = None) -> bool:
    """``window_attention`` on the tcgen05 kernels: CUDA fp16 / bf16 ``[B, Hi, Wi, P]`` projection (P >= 3 * heads * d,
    even), windows of at most 64 tokens that tile the token grid, even head dimension <= 64.  A bias that needs a gradient
    is served unless run-to-run determinism was asked for (its gradient is accumulated with float atomics)."""
    if not (qkv.is_cuda and qkv.dtype in _HALF and qkv.dim() == 4 and dispatch.enabled("attn") and dispatch._lib_ok()):
        return False
    if bias is not None and bias.requires_grad and dispatch.deterministic():
        return False
    _, Hi, Wi, P = qkv.shape
    return (qkv.is_contiguous() and P >= 3 * heads * d and P % 2 == 0 and 0 < ws * ws <= 64 and Hi % ws == 0
            and Wi % ws == 0 and 2 <= d <= 64 and d % 2 == 0)


class _WindowAttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, heads, d, ws, shift, scale, bias, mask, out_pitch):  # type: ignore[override]
        from .fused import DT
        lib = require_library()
        B, Hi, Wi, P = qkv.shape
        S = ws * ws
        nwin = B * (Hi // ws) * (Wi // ws)
        # columns past heads * d of a padded output row stay zero
        out = (torch.empty if out_pitch == heads * d else torch.zeros)((B, Hi, Wi, out_pitch), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((nwin, heads, S), dtype=torch.float32, device=qkv.device)
        b32 = None if bias is None else bias.detach().float().contiguous()        # [H, S, S]
        m32 = None if (mask is None or shift == 0) else mask.detach().float().contiguous()   # [windows per image, S, S]
        check(lib.b200dp_window_attention_fwd(ptr(qkv), ptr(out), ptr(lse), ptr(b32), ptr(m32), B, Hi, Wi, ws, shift, heads, d,
                                              P, out_pitch, ctypes.c_float(scale), DT[qkv.dtype], stream_ptr(qkv.device)),
              "window_attention_fwd")
        ctx.save_for_backward(qkv, out, lse, b32, m32)
        ctx.geom = (B, Hi, Wi, P, heads, d, ws, shift, scale, out_pitch)
        ctx.bias_dtype = None if bias is None else bias.dtype
        return out

    @staticmethod
    def backward(ctx, dout):  # type: ignore[override]
        from .fused import DT
        lib = require_library()
        qkv, out, lse, b32, m32 = ctx.saved_tensors
        B, Hi, Wi, P, H, d, ws, shift, scale, out_pitch = ctx.geom
        S = ws * ws
        if dout.dtype != qkv.dtype:
            dout = dout.to(qkv.dtype)
        if not dout.is_contiguous():
            dout = dout.contiguous()
        dqkv = (torch.empty_like if P == 3 * H * d else torch.zeros_like)(qkv)
        dbias = None
        if b32 is not None and ctx.needs_input_grad[6]:
            dbias = torch.zeros((H, S, S), dtype=torch.float32, device=qkv.device)
        check(lib.b200dp_window_attention_bwd(ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(dqkv), ptr(b32), ptr(m32), ptr(dbias),
                                              B, Hi, Wi, ws, shift, H, d, P, out_pitch, dout.shape[-1],
                                              ctypes.c_float(scale), DT[qkv.dtype], stream_ptr(qkv.device)),
              "window_attention_bwd")
        if dbias is not None and dbias.dtype != ctx.bias_dtype:
            dbias = dbias.to(ctx.bias_dtype)
        return dqkv, None, None, None, None, None, dbias, None, None


def window_attention(qkv: torch.Tensor, heads: int, d: int, ws: int, shift: int, scale: float,
                     bias: Optional[torch.Tensor] = None, mask: Optional[torch.Tensor] = None,
                     out_pitch: Optional[int] = None) -> torch.Tensor:
    """(Shifted-)window attention on the image-ordered token grid, no partition / roll copies.

    ``qkv [B, Hi, Wi, P]``: per token the packed projection (first ``3 * heads * d`` of its ``P`` elements, memory order
    ``[3][heads][d]``; ``P`` may be padded).  The grid is cyclically shifted by ``-shift`` in both directions, cut into
    ``ws x ws`` windows, ``softmax(q k^T * scale + bias + mask[window]) v`` runs per window and head, and the result goes
    back to each token's own position: ``[B, Hi, Wi, out_pitch]`` (first ``heads * d`` elements; the rest zeros).
    ``bias [heads, S, S]`` may require grad; ``mask [windows per image, S, S]`` is the shift mask (ignored when
    ``shift == 0``).  Equivalent to the roll -> window_partition -> attention -> window_reverse -> roll chain of a Swin
    block (Stoke-DDP.py:206-208 builds SwinIR with window_size 8).  Call ``window_attention_supported`` first."""
    return _WindowAttentionFn.apply(qkv, heads, d, ws, shift, float(scale), bias, mask,
                                    int(out_pitch) if out_pitch else heads * d)


def window_attention_reference(qkv: torch.Tensor, heads: int, d: int, ws: int, shift: int, scale: float,
                               bias: Optional[torch.Tensor] = None, mask: Optional[torch.Tensor] = None,
                               out_pitch: Optional[int] = None, dropout_p: float = 0.0) -> torch.Tensor:
    """Same contract with torch ops (CPU, shapes outside the envelope, attention dropout)."""
    import torch.nn.functional as F
    B, Hi, Wi, _ = qkv.shape
    C3 = 3 * heads * d
    x = qkv[..., :C3]
    if shift > 0:
        x = torch.roll(x, shifts=(-shift, -shift), dims=(1, 2))
    xw = x.reshape(B, Hi // ws, ws, Wi // ws, ws, C3).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, C3)
    m = mask if shift > 0 else None
    if attention_supported(xw, heads, bias) and dropout_p == 0.0:
        o = attention(xw.contiguous(), heads, scale, bias, m)
    else:
        o = attention_reference(xw, heads, scale, bias, m, dropout_p=dropout_p)
    C = heads * d
    o = o.reshape(B, Hi // ws, Wi // ws, ws, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(B, Hi, Wi, C)
    if shift > 0:
        o = torch.roll(o, shifts=(shift, shift), dims=(1, 2))
    if out_pitch and out_pitch != C:
        o = F.pad(o, (0, out_pitch - C))
    return o


def attention_reference(qkv: torch.Tensor, heads: int, scale: Optional[float] = None, bias: Optional[torch.Tensor] = None,
                        mask: Optional[torch.Tensor] = None, dropout_p: float = 0.0) -> torch.Tensor:
    """Same contract on the library kernel (shapes outside the envelope, CPU, attention dropout)."""
    import torch.nn.functional as F
    B, S, _ = qkv.shape
    q, k, v = split_qkv(qkv, heads)
    attn_mask = None
    if bias is not None:
        attn_mask = bias.unsqueeze(0).to(q.dtype)
        if mask is not None:
            nW = mask.shape[0]
            m = mask.to(q.dtype).view(1, nW, 1, S, S).expand(B // nW, -1, -1, -1, -1).reshape(B, 1, S, S)
            attn_mask = attn_mask + m
        else:
            attn_mask = attn_mask.expand(B, -1, -1, -1)
    elif mask is not None:
        nW = mask.shape[0]
        attn_mask = mask.to(q.dtype).view(1, nW, 1, S, S).expand(B // nW, -1, -1, -1, -1).reshape(B, 1, S, S)
    o = F.scaled_dot_product_attention(q, k, v, attn_mask=attn_mask, dropout_p=dropout_p, scale=scale)
    return merge_heads(o)


def split_qkv(qkv: torch.Tensor, heads: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """``[B, S, 3*D]`` projection -> q, k, v ``[B, H, S, d]`` (strided views)."""
    if not qkv.is_contiguous():
        qkv = qkv.contiguous()
    return _SplitQKV.apply(qkv, heads)


def merge_heads(o: torch.Tensor) -> torch.Tensor:
    """``[B, H, S, d]`` attention output -> ``[B, S, H*d]``."""
    return _MergeHeads.apply(o)

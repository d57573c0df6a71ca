"""Loss functions backed by ``csrc/loss.cu`` (W9).

Parity target: ``nn.MSELoss()`` (Fairscale-DDP.py:76); cross-entropy for the
classification models BASELINE.json names.  Both are plain callables usable as the
facade's ``loss=`` argument (Stoke-DDP.py:224,244).

The cross-entropy kernel reads the logits once and produces the mean loss *and* the
gradient ``(softmax - onehot)/rows`` in the same pass; backward only multiplies that
saved gradient by the upstream scalar (which carries the AMP loss scale).
"""
from __future__ import annotations

import ctypes

import torch
import torch.nn.functional as F

from . import dispatch
from ._lib import check, require_library
from .fused import DT, Workspace, ptr, stream_ptr

__all__ = ["cross_entropy", "mse_loss", "CrossEntropyLoss", "MSELoss"]


class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: torch.Tensor, labels: torch.Tensor, smoothing: float) -> torch.Tensor:  # type: ignore[override]
        lib = require_library()
        logits = logits.contiguous()
        rows, classes = logits.shape
        loss_rows = torch.zeros(rows + 1, dtype=torch.float32, device=logits.device)  # +1: completion ticket
        need_grad = ctx.needs_input_grad[0]
        dlogits = torch.empty_like(logits) if need_grad else None
        out = torch.empty((), dtype=torch.float32, device=logits.device)
        check(lib.b200dp_cross_entropy(ptr(logits), ptr(labels), ctypes.c_int(rows), ctypes.c_int(classes),
                                       ctypes.c_int(DT[logits.dtype]), ptr(loss_rows), ptr(dlogits), ptr(out),
                                       ctypes.c_float(smoothing), stream_ptr(logits.device)), "cross_entropy")
        ctx.dlogits = dlogits
        return out

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):  # type: ignore[override]
        lib = require_library()
        d = ctx.dlogits
        ctx.dlogits = None
        g = grad_out.detach().float().contiguous()
        check(lib.b200dp_scale_inplace(ptr(d), ctypes.c_int64(d.numel()), ctypes.c_int(DT[d.dtype]), ptr(g),
                                       stream_ptr(d.device)), "scale_inplace")
        return d, None, None


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor, label_smoothing: float = 0.0) -> torch.Tensor:
    """Mean softmax cross-entropy over ``logits[rows, classes]`` with int64 ``labels``.
    Returns an fp32 scalar."""
    if (logits.is_cuda and logits.dim() == 2 and logits.dtype in DT and labels.dtype == torch.int64
            and dispatch.enabled("ce") and dispatch._lib_ok()):
        return _CrossEntropyFn.apply(logits, labels.contiguous(), float(label_smoothing))
    return F.cross_entropy(logits.float(), labels, label_smoothing=label_smoothing)


class _MSEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:  # type: ignore[override]
        lib = require_library()
        a, b = a.contiguous(), b.contiguous()
        n = a.numel()
        ws = Workspace.get(a.device)
        scratch = ws.scratch(int(lib.b200dp_sumsq_scratch_floats(ctypes.c_int64(n))))
        grad = torch.empty_like(a) if (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]) else None
        out = torch.empty((), dtype=torch.float32, device=a.device)
        check(lib.b200dp_mse(ptr(a), ptr(b), ctypes.c_int64(n), ctypes.c_int(DT[a.dtype]), ptr(grad), ptr(scratch),
                             ptr(ws.counters), ptr(out), stream_ptr(a.device)), "mse")
        ctx.grad = grad
        return out

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):  # type: ignore[override]
        lib = require_library()
        d = ctx.grad
        ctx.grad = None
        g = grad_out.detach().float().contiguous()
        check(lib.b200dp_scale_inplace(ptr(d), ctypes.c_int64(d.numel()), ctypes.c_int(DT[d.dtype]), ptr(g),
                                       stream_ptr(d.device)), "scale_inplace")
        da = d if ctx.needs_input_grad[0] else None
        db = -d if ctx.needs_input_grad[1] else None
        return da, db


def mse_loss(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    if (a.is_cuda and a.dtype in DT and b.dtype == a.dtype and a.shape == b.shape and dispatch.enabled("mse")
            and dispatch._lib_ok()):
        return _MSEFn.apply(a, b)
    return F.mse_loss(a.float(), b.float())


class CrossEntropyLoss(torch.nn.Module):
    def __init__(self, label_smoothing: float = 0.0):
        super().__init__()
        self.label_smoothing = label_smoothing

    def forward(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        return cross_entropy(logits, labels, self.label_smoothing)


class MSELoss(torch.nn.Module):
    def forward(self, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
        return mse_loss(a, b)

"""BatchNorm(+residual+ReLU) and LayerNorm autograd functions over ``csrc/norm.cu``
(W8; the local half of W6 — cross-rank statistics are exchanged between the reduce
and the apply kernel).

Parity targets: the BatchNorm layers of the CNN (Fairscale-DDP.py:74), the
``convert_to_sync_batch_norm=True`` flag (Stoke-DDP.py:192) and the LayerNorms of the
transformer models (Stoke-DDP.py:206-208).
"""
from __future__ import annotations

import ctypes
from typing import Any, Optional

import torch
import torch.distributed as dist

from ._lib import check, require_library
from .fused import DT, Workspace, ptr, stream_ptr


def _rows_view(x: torch.Tensor):
    """NCHW-shaped channels-last tensor -> (rows, C)."""
    n, c, h, w = x.shape
    return n * h * w, c


def _allreduce_stats(buf: torch.Tensor, group: Any) -> None:
    """Cross-rank sum of the moment vector (K6/K7): peer-memory one-shot all-reduce when
    attached, NCCL otherwise."""
    from ..parallel import peer
    if not peer.small_allreduce(buf, group):
        dist.all_reduce(buf, group=group)


def column_sums(x2: torch.Tensor) -> torch.Tensor:
    """fp32 column sums of a contiguous 16-bit ``[rows, C]`` matrix (one streaming read plus a small second stage;
    used for bias gradients).  ``C % 8 == 0`` and ``C <= 16384``."""
    lib = require_library()
    rows, C = x2.shape
    ws = Workspace.get(x2.device)
    scratch = ws.scratch(int(lib.b200dp_column_sums_scratch_floats(ctypes.c_int64(rows), ctypes.c_int(C))))
    sums = torch.empty(C, dtype=torch.float32, device=x2.device)
    check(lib.b200dp_column_sums(ptr(x2), ctypes.c_int64(rows), ctypes.c_int(C), ctypes.c_int(DT[x2.dtype]), ptr(scratch),
                                 ptr(sums), stream_ptr(x2.device)), "column_sums", launches=2)
    return sums


class _BatchNormActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, running_mean, running_var, eps, momentum, relu, group, world,
                pre_sums=None, tracked=None):  # type: ignore[override]
        lib = require_library()
        rows, C = _rows_view(x)
        dt = ctypes.c_int(DT[x.dtype])
        dev = x.device
        ws = Workspace.get(dev)
        st = stream_ptr(dev)
        if pre_sums is not None:
            sums = pre_sums            # produced by the epilogue of the convolution that wrote x
        else:
            scratch = ws.scratch(int(lib.b200dp_bn_scratch_floats(ctypes.c_int64(rows), ctypes.c_int(C))))
            sums = torch.empty(2 * C + 1, dtype=torch.float32, device=dev)
            check(lib.b200dp_bn_reduce(ptr(x), ctypes.c_int64(rows), ctypes.c_int(C), dt, ptr(scratch), ptr(ws.counters),
                                       ptr(sums), st), "bn_reduce")
        count = float(rows)
        if world > 1:
            sums[2 * C: 2 * C + 1].fill_(count)          # (a fill kernel: indexed assignment would be a host copy, illegal under capture)
            _allreduce_stats(sums, group)
            # the kernels read the TRUE element count from slot 2C of the reduced vector (count <= 0 selects that):
            # no host sync, and a ragged last batch on one rank (validation, drop_last=False) does not skew the statistics
            count = 0.0
        y = torch.empty_like(x)
        mean = torch.empty(C, dtype=torch.float32, device=dev)
        invstd = torch.empty(C, dtype=torch.float32, device=dev)
        mask = torch.empty(rows * (C // 8), dtype=torch.uint8, device=dev) if relu else None   # 1 bit / element
        # `tracked`: the module's int64 num_batches_tracked, incremented by one thread of the kernel
        check(lib.b200dp_bn_apply_tracked(ptr(x), ptr(residual), ptr(y), ctypes.c_int64(rows), ctypes.c_int(C), dt, ptr(sums),
                                          ctypes.c_float(count), ptr(weight), ptr(bias), ctypes.c_float(eps),
                                          ctypes.c_float(momentum), ptr(running_mean), ptr(running_var), ptr(mean),
                                          ptr(invstd), ctypes.c_int(int(relu)), ptr(mask), ptr(tracked), st), "bn_apply")
        ctx.save_for_backward(x, mask, weight, mean, invstd)
        ctx.relu, ctx.group, ctx.world, ctx.count = relu, group, world, count
        ctx.total = sums[2 * C: 2 * C + 1] if world > 1 else None      # device scalar: cross-rank element count
        ctx.has_res = residual is not None
        return y

    @staticmethod
    def backward(ctx, dy):  # type: ignore[override]
        lib = require_library()
        x, mask, weight, mean, invstd = ctx.saved_tensors
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        rows, C = _rows_view(x)
        dt = ctypes.c_int(DT[x.dtype])
        dev = x.device
        ws = Workspace.get(dev)
        scratch = ws.scratch(int(lib.b200dp_bn_scratch_floats(ctypes.c_int64(rows), ctypes.c_int(C))))
        dsums = torch.empty(2 * C + 1, dtype=torch.float32, device=dev)     # [sum dz | sum dz*xhat | element count]
        st = stream_ptr(dev)
        relu = ctypes.c_int(int(ctx.relu))
        check(lib.b200dp_bn_bwd_reduce(ptr(dy), None, ptr(x), ctypes.c_int64(rows), ctypes.c_int(C), dt, ptr(mean),
                                       ptr(invstd), relu, ptr(scratch), ptr(ws.counters), ptr(dsums), ptr(mask), st),
              "bn_bwd_reduce")
        # parameter gradients are the LOCAL sums (the data-parallel engine averages them); they only need
        # their own storage when the cross-rank exchange is about to overwrite dsums
        dbias, dweight = dsums[:C], dsums[C: 2 * C]
        if ctx.world > 1:
            dbias, dweight = dbias.clone(), dweight.clone()
            _allreduce_stats(dsums[: 2 * C], ctx.group)
            dsums[2 * C: 2 * C + 1].copy_(ctx.total)
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if ctx.has_res else None
        check(lib.b200dp_bn_bwd_apply(ptr(dy), None, ptr(x), ptr(dx), ptr(dres), ctypes.c_int64(rows), ctypes.c_int(C),
                                      dt, ptr(mean), ptr(invstd), ptr(weight), ptr(dsums), ctypes.c_float(ctx.count),
                                      relu, ptr(mask), st), "bn_bwd_apply")
        return dx, dres, dweight.to(weight.dtype), dbias.to(weight.dtype), None, None, None, None, None, None, None, None, None


def batch_norm_act(mod: Any, x: torch.Tensor, residual: Optional[torch.Tensor]) -> torch.Tensor:
    """``act(bn(x) + residual)`` for a ``BatchNormAct2d`` module on NHWC 16-bit input."""
    lib = require_library()
    relu = mod.act == "relu"
    if residual is not None and not residual.is_contiguous(memory_format=torch.channels_last):
        residual = residual.contiguous(memory_format=torch.channels_last)
    if residual is not None and residual.dtype != x.dtype:
        residual = residual.to(x.dtype)
    if not mod.training:
        rows, C = _rows_view(x)
        y = torch.empty_like(x)
        check(lib.b200dp_bn_infer(ptr(x), ptr(residual), ptr(y), ctypes.c_int64(rows), ctypes.c_int(C),
                                  ctypes.c_int(DT[x.dtype]), ptr(mod.weight), ptr(mod.bias), ptr(mod.running_mean),
                                  ptr(mod.running_var), ctypes.c_float(mod.eps), ctypes.c_int(int(relu)),
                                  stream_ptr(x.device)), "bn_infer")
        return y
    momentum = 0.1 if mod.momentum is None else mod.momentum
    nbt = mod.num_batches_tracked
    if nbt is not None and not (nbt.is_cuda and nbt.dtype == torch.int64 and nbt.device == x.device):
        nbt.add_(1)                 # not where the kernel can reach it
        nbt = None
    world, group = 1, None
    if mod.sync and dist.is_available() and dist.is_initialized():
        group = mod.process_group if mod.process_group is not None else dist.group.WORLD
        world = dist.get_world_size(group)
    pre = getattr(x, "_b200dp_bn_sums", None)
    if pre is not None and (pre.numel() != 2 * x.shape[1] + 1 or pre.device != x.device):
        pre = None
    return _BatchNormActFn.apply(x, residual, mod.weight, mod.bias, mod.running_mean, mod.running_var, mod.eps,
                                 momentum, relu, group, world, pre, nbt)


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, carry=False, logical=0):  # type: ignore[override]
        """``logical``: when non-zero, rows of ``x`` are ``x.shape[-1]`` elements apart but only the first
        ``logical`` of them are normalised (zero-padded rows, e.g. 60 -> 64); the padding of ``y`` / ``dx`` is zeros."""
        lib = require_library()
        # parameters arrive in their own dtype (bf16 under FSDP mixed precision); the kernels read fp32.  Casting
        # here instead of in the caller keeps autograd from adding one cast-backward kernel per parameter.
        ctx.p_dtype = weight.dtype
        weight, bias = weight.float(), bias.float()
        shape = x.shape
        D = shape[-1]
        x2 = x.reshape(-1, D)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        ctx.logical = int(logical) if logical and int(logical) != D else 0
        if ctx.logical:
            check(lib.b200dp_layernorm_fwd_pitched(ptr(x2), ptr(y), ctypes.c_int64(rows), ctypes.c_int(ctx.logical),
                                                   ctypes.c_int(D), ctypes.c_int(DT[x.dtype]), ptr(weight), ptr(bias),
                                                   ctypes.c_float(eps), ptr(mean), ptr(rstd), stream_ptr(x.device)),
                  "layernorm_fwd_pitched")
        else:
            check(lib.b200dp_layernorm_fwd(ptr(x2), ptr(y), ctypes.c_int64(rows), ctypes.c_int(D), ctypes.c_int(DT[x.dtype]),
                                           ptr(weight), ptr(bias), ctypes.c_float(eps), ptr(mean), ptr(rstd),
                                           stream_ptr(x.device)), "layernorm_fwd")
        ctx.save_for_backward(x2, weight, mean, rstd)
        ctx.shape = shape
        if carry:
            # second output: the input itself, for the residual branch that bypasses the block; its gradient
            # comes back into this backward and is added to dx inside the kernel
            return y.view(shape), x.view_as(x)
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy, d_carry=None):  # type: ignore[override]
        lib = require_library()
        x2, weight, mean, rstd = ctx.saved_tensors
        rows, D = x2.shape
        dy2 = dy.reshape(rows, D)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        if dy2.dtype != x2.dtype:
            dy2 = dy2.to(x2.dtype)
        add_in = None
        if d_carry is not None:
            add_in = d_carry.reshape(rows, D)
            if add_in.dtype != x2.dtype:
                add_in = add_in.to(x2.dtype)
            if not add_in.is_contiguous():
                add_in = add_in.contiguous()
        dx = torch.empty_like(x2)
        ws = Workspace.get(x2.device)
        if ctx.logical:
            P, D = D, ctx.logical
            scratch = ws.scratch(int(lib.b200dp_ln_scratch_floats(ctypes.c_int64(rows), ctypes.c_int(D))))
            dparams = torch.empty(2 * D, dtype=torch.float32, device=x2.device)  # [dbeta | dgamma]
            check(lib.b200dp_layernorm_bwd_pitched(ptr(dy2), ptr(x2), ptr(dx), ctypes.c_int64(rows), ctypes.c_int(D),
                                                   ctypes.c_int(P), ctypes.c_int(DT[x2.dtype]), ptr(weight), ptr(mean),
                                                   ptr(rstd), ptr(scratch), ptr(ws.counters), ptr(dparams[D:]),
                                                   ptr(dparams[:D]), ptr(add_in), stream_ptr(x2.device)),
                  "layernorm_bwd_pitched", launches=2)
            if dparams.dtype != ctx.p_dtype:
                dparams = dparams.to(ctx.p_dtype)
            return dx.view(ctx.shape), dparams[D:], dparams[:D], None, None, None
        scratch = ws.scratch(int(lib.b200dp_ln_scratch_floats(ctypes.c_int64(rows), ctypes.c_int(D))))
        dparams = torch.empty(2 * D, dtype=torch.float32, device=x2.device)  # [dbeta | dgamma]
        check(lib.b200dp_layernorm_bwd_add(ptr(dy2), ptr(x2), ptr(dx), ctypes.c_int64(rows), ctypes.c_int(D),
                                           ctypes.c_int(DT[x2.dtype]), ptr(weight), ptr(mean), ptr(rstd), ptr(scratch),
                                           ptr(ws.counters), ptr(dparams[D:]), ptr(dparams[:D]), ptr(add_in),
                                           stream_ptr(x2.device)),
              "layernorm_bwd", launches=2)
        if dparams.dtype != ctx.p_dtype:
            dparams = dparams.to(ctx.p_dtype)           # one cast for both parameter gradients
        return dx.view(ctx.shape), dparams[D:], dparams[:D], None, None, None


def layer_norm(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, eps: float, carry: bool = False,
               logical: int = 0):
    """``carry=True`` returns ``(LN(x), x)``: use the second value as the residual that bypasses the block.
    ``logical``: normalise only the first ``logical`` elements of zero-padded rows (see ``_LayerNormFn.forward``)."""
    return _LayerNormFn.apply(x, weight, bias, eps, carry, logical)

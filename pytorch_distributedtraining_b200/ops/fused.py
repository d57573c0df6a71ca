"""ctypes bindings for the multi-tensor / AMP / norm kernels of ``csrc/optim.cu``.

Everything here launches on the current torch CUDA stream and never synchronises.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ._lib import check, require_library

CHUNK = 8192
DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
_DESC = np.dtype([("p", "<u8"), ("g", "<u8"), ("m", "<u8"), ("v", "<u8"), ("lp", "<u8"), ("n", "<i8")])


class AdamHParams(ctypes.Structure):
    _fields_ = [("lr", ctypes.c_float), ("beta1", ctypes.c_float), ("beta2", ctypes.c_float),
                ("eps", ctypes.c_float), ("weight_decay", ctypes.c_float), ("adamw", ctypes.c_int),
                ("maximize", ctypes.c_int), ("lr_dev", ctypes.c_void_p)]


class SGDHParams(ctypes.Structure):
    _fields_ = [("lr", ctypes.c_float), ("momentum", ctypes.c_float), ("dampening", ctypes.c_float),
                ("weight_decay", ctypes.c_float), ("nesterov", ctypes.c_int), ("maximize", ctypes.c_int),
                ("lr_dev", ctypes.c_void_p)]


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr(device: Optional[torch.device] = None) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


# --------------------------------------------------------------------------- workspace
class Workspace:
    """Per-(device, stream) scratch: a growable float buffer plus 64 zeroed completion
    counters.  Kernels leave the counters zero, so they are allocated and cleared once."""

    _cache: Dict[Tuple[int, int], "Workspace"] = {}

    def __init__(self, device: torch.device):
        self.device = device
        self.counters = torch.zeros(64, dtype=torch.int32, device=device)
        self._scratch = torch.empty(1 << 16, dtype=torch.float32, device=device)

    @classmethod
    def get(cls, device: torch.device) -> "Workspace":
        key = (device.index if device.index is not None else torch.cuda.current_device(),
               torch.cuda.current_stream(device).cuda_stream)
        ws = cls._cache.get(key)
        if ws is None:
            ws = cls._cache[key] = Workspace(device)
        return ws

    def scratch(self, floats: int) -> torch.Tensor:
        if self._scratch.numel() < floats:
            self._scratch = torch.empty(int(floats * 1.5), dtype=torch.float32, device=self.device)
        return self._scratch


# --------------------------------------------------------------------------- tensor lists
def is_dense(t: torch.Tensor) -> bool:
    """Covers its storage span exactly once: plain contiguous or channels-last contiguous."""
    if t.is_contiguous():
        return True
    if t.dim() == 4:
        return t.is_contiguous(memory_format=torch.channels_last)
    if t.dim() == 5:
        return t.is_contiguous(memory_format=torch.channels_last_3d)
    return False


def same_layout(a: torch.Tensor, b: torch.Tensor) -> bool:
    """Same logical->physical element order (the kernels walk raw memory linearly)."""
    if a.shape != b.shape:
        return False
    if a.stride() == b.stride() or (a.is_contiguous() and b.is_contiguous()):
        return True
    if a.dim() == 4:
        return a.is_contiguous(memory_format=torch.channels_last) and b.is_contiguous(memory_format=torch.channels_last)
    return False


def dense_view_like(flat: torch.Tensor, p: torch.Tensor) -> torch.Tensor:
    """View of ``flat`` (``p.numel()`` elements) with ``p``'s shape AND memory layout."""
    if p.is_contiguous() or not is_dense(p):
        return flat.view(p.shape)
    return flat.as_strided(p.shape, p.stride())


class TensorList:
    """Device-side descriptor + chunk tables for a list of (param, grad, m, v, lp) tensors."""

    def __init__(self, p: Sequence[Optional[torch.Tensor]], g: Sequence[Optional[torch.Tensor]],
                 m: Optional[Sequence[Optional[torch.Tensor]]] = None,
                 v: Optional[Sequence[Optional[torch.Tensor]]] = None,
                 lp: Optional[Sequence[Optional[torch.Tensor]]] = None):
        n = len(g)
        if torch.cuda.is_available() and torch.cuda.is_current_stream_capturing():
            raise RuntimeError("a fused-kernel tensor table would have to be (re)built during CUDA-graph capture: "
                               "parameter / gradient addresses changed since the warm-up steps (keep gradients "
                               "allocated, e.g. zero_grad(set_to_none=False))")
        ref = next(t for t in list(g) + list(p) if t is not None)
        self.device = ref.device
        desc = np.zeros(n, dtype=_DESC)
        chunks: List[Tuple[int, int]] = []
        for i in range(n):
            t = g[i] if g[i] is not None else p[i]
            numel = t.numel()
            for name, seq in (("p", p), ("g", g), ("m", m), ("v", v), ("lp", lp)):
                x = seq[i] if seq is not None and i < len(seq) else None
                if x is not None:
                    if not is_dense(x):
                        raise ValueError("fused kernels need dense (contiguous or channels-last) tensors")
                    if not same_layout(x, t):
                        raise ValueError(f"fused kernels walk raw memory: '{name}' of tensor {i} has strides "
                                         f"{x.stride()} but its gradient has {t.stride()} (shape {tuple(t.shape)})")
                    desc[name][i] = x.data_ptr()
            desc["n"][i] = numel
            chunks.extend((i, c) for c in range((numel + CHUNK - 1) // CHUNK))
        self.n_chunks = len(chunks)
        self.key = tuple(int(x) for x in desc["g"]) + tuple(int(x) for x in desc["p"])
        self.desc = torch.from_numpy(desc.view(np.uint8).copy()).to(self.device)
        self.chunks = torch.tensor(chunks if chunks else [(0, 0)], dtype=torch.int32).to(self.device)
        self.keepalive = (p, g, m, v, lp)


def unscale_check_(grads: List[torch.Tensor], inv_scale: torch.Tensor, found_inf: torch.Tensor) -> None:
    """In-place ``g *= inv_scale`` over ``grads`` with a non-finite check (K13)."""
    lib = require_library()
    by_dtype: Dict[torch.dtype, List[torch.Tensor]] = {}
    for g in grads:
        by_dtype.setdefault(g.dtype, []).append(g)
    for dtype, gs in by_dtype.items():
        # tables are cached by the gradient addresses: steady-state training (and a CUDA-graph capture after the warm-up
        # steps) never rebuilds them
        key = (dtype,) + tuple(g.data_ptr() for g in gs)
        tl = _UNSCALE_TABLES.get(key)
        if tl is None:
            if len(_UNSCALE_TABLES) > 16:
                _UNSCALE_TABLES.clear()
            tl = _UNSCALE_TABLES[key] = TensorList([None] * len(gs), gs)
        check(lib.b200dp_multi_unscale_check(ptr(tl.desc), ptr(tl.chunks), ctypes.c_int(tl.n_chunks),
                                             ctypes.c_int(DT[dtype]), ptr(inv_scale), ptr(found_inf),
                                             stream_ptr(tl.device)), "multi_unscale_check")


_UNSCALE_TABLES: Dict[tuple, "TensorList"] = {}


_KEEP: List[object] = []


def _keep(obj: object, limit: int = 64) -> None:
    """Keep short-lived device tables alive until the stream has certainly consumed them
    (the caching allocator would otherwise be free to hand the memory out again)."""
    _KEEP.append(obj)
    if len(_KEEP) > limit:
        del _KEEP[: limit // 2]


def scaler_update_(scale: torch.Tensor, growth_tracker: torch.Tensor, found_inf: torch.Tensor, growth_factor: float,
                   backoff_factor: float, growth_interval: int) -> None:
    lib = require_library()
    check(lib.b200dp_scaler_update(ptr(scale), ptr(growth_tracker), ptr(found_inf), ctypes.c_float(growth_factor),
                                   ctypes.c_float(backoff_factor), ctypes.c_int(growth_interval),
                                   stream_ptr(scale.device)), "scaler_update")


def sumsq_check(tl: TensorList, grad_dtype: torch.dtype, out: torch.Tensor) -> None:
    """``out[0] = sum g^2`` over the list, ``out[1] = 1`` if anything is non-finite (K8)."""
    lib = require_library()
    ws = Workspace.get(tl.device)
    scratch = ws.scratch(max(tl.n_chunks, 1))
    check(lib.b200dp_multi_sumsq_check(ptr(tl.desc), ptr(tl.chunks), ctypes.c_int(tl.n_chunks),
                                       ctypes.c_int(DT[grad_dtype]), ptr(scratch), ptr(ws.counters), ptr(out),
                                       stream_ptr(tl.device)), "multi_sumsq_check")


def clip_coef(sumsq: torch.Tensor, inv_scale: Optional[torch.Tensor], max_norm: float, grad_mult: torch.Tensor,
              total_norm: Optional[torch.Tensor]) -> None:
    lib = require_library()
    check(lib.b200dp_clip_coef(ptr(sumsq), ptr(inv_scale), ctypes.c_float(max_norm), ptr(grad_mult),
                               ptr(total_norm), stream_ptr(sumsq.device)), "clip_coef")


def step_advance(step: torch.Tensor, found_inf: Optional[torch.Tensor]) -> None:
    lib = require_library()
    check(lib.b200dp_step_advance(ptr(step), ptr(found_inf), stream_ptr(step.device)), "step_advance")


def sumsq_f32(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Deterministic two-stage sum of squares of a flat fp32 tensor."""
    lib = require_library()
    lib.b200dp_sumsq_scratch_floats.restype = ctypes.c_int64
    assert x.dtype == torch.float32 and x.is_contiguous()
    ws = Workspace.get(x.device)
    n = x.numel()
    scratch = ws.scratch(int(lib.b200dp_sumsq_scratch_floats(ctypes.c_int64(n))))
    if out is None:
        out = torch.empty((), dtype=torch.float32, device=x.device)
    check(lib.b200dp_sumsq_f32(ptr(x), ctypes.c_int64(n), ptr(scratch), ptr(ws.counters), ptr(out),
                               stream_ptr(x.device)), "sumsq_f32")
    return out


def multi_adam(tl: TensorList, grad_dtype: torch.dtype, lp_dtype: torch.dtype, hp: AdamHParams, step: torch.Tensor,
               grad_mult: Optional[torch.Tensor], found_inf: Optional[torch.Tensor]) -> None:
    lib = require_library()
    check(lib.b200dp_multi_adam(ptr(tl.desc), ptr(tl.chunks), ctypes.c_int(tl.n_chunks), ctypes.c_int(DT[grad_dtype]),
                                ctypes.c_int(DT.get(lp_dtype, 1)), hp, ptr(step), ptr(grad_mult), ptr(found_inf),
                                stream_ptr(tl.device)), "multi_adam")


def multi_sgd(tl: TensorList, grad_dtype: torch.dtype, lp_dtype: torch.dtype, hp: SGDHParams, first_step: bool,
              grad_mult: Optional[torch.Tensor], found_inf: Optional[torch.Tensor]) -> None:
    lib = require_library()
    check(lib.b200dp_multi_sgd(ptr(tl.desc), ptr(tl.chunks), ctypes.c_int(tl.n_chunks), ctypes.c_int(DT[grad_dtype]),
                               ctypes.c_int(DT.get(lp_dtype, 1)), hp, ctypes.c_int(int(first_step)), ptr(grad_mult),
                               ptr(found_inf), stream_ptr(tl.device)), "multi_sgd")


# --------------------------------------------------------------------------- 16-bit weight shadows
LP_ATTR, LP_VER = "_b200dp_lp", "_b200dp_lp_version"


def lp_weight(weight: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """The 16-bit copy of an fp32 weight for a tensor-core kernel.  When the optimizer keeps a shadow of the parameter in
    ``dtype`` (written by the SAME kernel that updates the fp32 master, or the 16-bit wire copy a ZeRO owner published)
    and the parameter has not been touched since, that shadow is the operand — no per-forward cast kernel; otherwise
    one ``.to(dtype)``."""
    if weight.dtype == dtype:
        return weight
    lp = getattr(weight, LP_ATTR, None)
    if (lp is not None and lp.dtype == dtype and lp.shape == weight.shape
            and getattr(weight, LP_VER, -1) == weight._version):
        return lp
    return weight.to(dtype)


# Bumped whenever parameters are updated by a path that does not touch their version counters: the fused optimizer
# kernels write through raw pointers, and a CUDA-graph replay runs no Python at all.  Caches derived from parameter
# VALUES (``padded_param``) key on (parameter version, this epoch).
PARAM_EPOCH = 0


def bump_param_epoch() -> None:
    global PARAM_EPOCH
    PARAM_EPOCH += 1


class _PaddedParam(torch.autograd.Function):
    """A parameter zero-padded to ``shape`` in a persistent buffer, refreshed only when the parameter changed (once per
    optimizer step — not once per micro-batch and not once per use), together with its 16-bit twin for the tensor-core
    kernels (hung on the result the way ``lp_weight`` expects, so the consumer runs no cast kernel).  Backward: the slice
    of the incoming gradient that belongs to the parameter."""

    @staticmethod
    def forward(ctx, p, shape, holder, lp_dtype):  # type: ignore[override]
        ctx.core = tuple(slice(0, n) for n in p.shape)
        buf = holder.get("buf")
        if buf is None or buf.shape != torch.Size(shape) or buf.device != p.device or buf.dtype != p.dtype:
            buf = holder["buf"] = torch.zeros(shape, dtype=p.dtype, device=p.device)
            holder["key"], holder["lp"] = None, None
        key = (p._version, PARAM_EPOCH, p.data_ptr())
        fresh = holder.get("key") != key
        if fresh:
            buf[ctx.core].copy_(p.detach())
            holder["key"] = key
        out = buf.view(shape)               # a new tensor object per call (the buffer itself never gets a grad_fn)
        if lp_dtype is not None and lp_dtype != p.dtype:
            lp = holder.get("lp")
            if lp is None or lp.dtype != lp_dtype or lp.shape != buf.shape:
                # convolution filters: the kernels want the 16-bit copy channels-last ([Cout][R][S][Cin] memory)
                mf = torch.channels_last if buf.dim() == 4 else torch.contiguous_format
                lp = holder["lp"] = torch.empty_like(buf, dtype=lp_dtype, memory_format=mf)
                fresh = True
            if fresh:
                lp.copy_(buf)
            setattr(out, LP_ATTR, lp)
            setattr(out, LP_VER, out._version)
        return out

    @staticmethod
    def backward(ctx, g):  # type: ignore[override]
        return g[ctx.core], None, None, None


def padded_param(p: torch.Tensor, shape, holder: dict, lp_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """``p`` zero-padded to ``shape`` (see ``_PaddedParam``); ``holder`` is the caller's cache slot (one dict per
    parameter, e.g. on the owning module).  Parameters that are re-materialised every forward (FSDP views) skip the
    cache: their identity, not just their value, changes."""
    if tuple(shape) == tuple(p.shape):
        return p
    if not isinstance(p, torch.nn.Parameter):
        pad = []
        for have, want in zip(reversed(p.shape), reversed(tuple(shape))):
            pad += [0, want - have]
        return torch.nn.functional.pad(p, pad)
    return _PaddedParam.apply(p, tuple(shape), holder, lp_dtype)


def stamp_lp(params) -> None:
    for p in params:
        if getattr(p, LP_ATTR, None) is not None:
            setattr(p, LP_VER, p._version)

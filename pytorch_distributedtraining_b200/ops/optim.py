"""Fused sm_100a optimizers with stock ``torch.optim`` state_dict layouts (W10 / K14).

Parity target: ``AdamW(lr=1e-3, betas=(0.9, 0.99), eps=1e-8, weight_decay=...)``
(Fairscale-DDP.py:78-84, Stoke-DDP.py:226-235); ``SGD`` is imported by the reference
(Stoke-DDP.py:27) and is the default of OSS.

One kernel launch updates every parameter of a param group (multi-tensor chunk table,
``csrc/optim.cu``).  Fused in: the AMP unscale, the clipping coefficient (both arrive
as one device scalar ``grad_mult``) and the GradScaler's skip-on-inf decision (a device
flag, so ``step`` never reads anything back to the host).  State tensors are views into
flat per-group buffers; ``state_dict()`` is exactly torch's
(``step`` / ``exp_avg`` / ``exp_avg_sq`` / ``momentum_buffer``).
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Tuple, Type

import torch
from torch.optim import Optimizer

from . import fused

__all__ = ["FusedAdamW", "FusedAdam", "FusedSGD", "fused_equivalent"]


def _flat_like(params: List[torch.Tensor]) -> Tuple[torch.Tensor, List[torch.Tensor]]:
    """One zeroed fp32 buffer with a 16-byte aligned view per parameter."""
    offs, total = [], 0
    for p in params:
        offs.append(total)
        total += (p.numel() + 3) // 4 * 4
    flat = torch.zeros(max(total, 4), dtype=torch.float32, device=params[0].device)
    return flat, [fused.dense_view_like(flat[o: o + p.numel()], p) for o, p in zip(offs, params)]


class _FusedBase(Optimizer):
    supports_device_skip = True

    def __init__(self, params, defaults):
        super().__init__(params, defaults)
        self._tables: Dict[int, fused.TensorList] = {}
        self._grad_mult: Optional[torch.Tensor] = None   # set by fused_unscale_clip, consumed by step
        self._norm_buf: Optional[torch.Tensor] = None
        self._total_norm: Optional[torch.Tensor] = None
        self._norm_table: Optional[fused.TensorList] = None

    # 16-bit shadows of the matrix-shaped parameters, written by the update kernel itself (``desc.p_lp``): the
    # convolution / linear kernels read them instead of casting the fp32 weights every forward.  Set by the facade to
    # the autocast dtype; None keeps the optimizer shadow-free.
    lp_dtype: Optional[torch.dtype] = None

    def _shadows(self, params: List[torch.Tensor]) -> Optional[List[Optional[torch.Tensor]]]:
        if self.lp_dtype is None:
            return None
        out: List[Optional[torch.Tensor]] = []
        for p in params:
            if p.dim() < 2:
                out.append(None)
                continue
            lp = getattr(p, fused.LP_ATTR, None)
            if lp is None or lp.dtype != self.lp_dtype or lp.shape != p.shape or not fused.same_layout(lp, p):
                lp = torch.empty_like(p, dtype=self.lp_dtype)
                lp.copy_(p.detach())
                setattr(p, fused.LP_ATTR, lp)
                setattr(p, fused.LP_VER, p._version)
            out.append(lp)
        return out

    def make_shadows(self) -> None:
        """Create the shadows up front (before the first step), so the very first forward already uses them."""
        for g in self.param_groups:
            self._shadows([p for p in g["params"] if p.requires_grad])

    # -- helpers ---------------------------------------------------------------
    def _params_with_grad(self, group) -> List[torch.Tensor]:
        out = []
        for p in group["params"]:
            if p.grad is not None:
                if p.grad.is_sparse:
                    raise RuntimeError("fused optimizers do not support sparse gradients")
                if p.dtype != torch.float32:
                    raise RuntimeError("fused optimizers expect fp32 master parameters")
                out.append(p)
        return out

    def _all_grads(self) -> List[torch.Tensor]:
        return [p.grad for g in self.param_groups for p in g["params"] if p.grad is not None]

    def _device(self) -> torch.device:
        return self.param_groups[0]["params"][0].device

    @torch.no_grad()
    def fused_unscale_clip(self, inv_scale: Optional[torch.Tensor], max_norm: Optional[float],
                           found_inf: Optional[torch.Tensor], extra_sumsq_reduce: Optional[Callable] = None
                           ) -> Optional[torch.Tensor]:
        """ONE read of all gradients: sum of squares + non-finite flag; then a one-thread
        kernel turns (sumsq, 1/scale, max_norm) into the multiplier the update kernel
        applies.  Nothing is written back to the gradients.

        ``extra_sumsq_reduce(buf)``: hook for sharded optimizers to all-reduce the partial
        sum of squares (and flag) across ranks before the coefficient is computed.
        Returns the total gradient norm (device scalar) when clipping, else ``None``.
        """
        grads = self._all_grads()
        if not grads:
            return None
        dev = grads[0].device
        if self._norm_buf is None:
            self._norm_buf = torch.zeros(2, dtype=torch.float32, device=dev)
            self._grad_mult = torch.ones((), dtype=torch.float32, device=dev)
            self._total_norm = torch.zeros((), dtype=torch.float32, device=dev)
        key = tuple(g.data_ptr() for g in grads)
        if self._norm_table is None or self._norm_table.key[: len(key)] != key:
            self._norm_table = fused.TensorList([None] * len(grads), grads)
        self._norm_buf.zero_()
        fused.sumsq_check(self._norm_table, grads[0].dtype, self._norm_buf)
        if extra_sumsq_reduce is not None:
            extra_sumsq_reduce(self._norm_buf)
        if found_inf is not None:
            torch.maximum(found_inf, self._norm_buf[1], out=found_inf)
        fused.clip_coef(self._norm_buf[0:1], inv_scale, float(max_norm) if max_norm is not None else 3.0e38,
                        self._grad_mult, self._total_norm)
        self._use_grad_mult = True
        return self._total_norm if max_norm is not None else None

    def clip_grad_norm(self, max_norm: float, norm_type: float = 2.0) -> torch.Tensor:
        if float(norm_type) != 2.0:
            total = torch.nn.utils.clip_grad_norm_(
                [p for g in self.param_groups for p in g["params"] if p.grad is not None], max_norm, norm_type)
            return total
        return self.fused_unscale_clip(None, max_norm, None)

    @staticmethod
    def _bump_versions(params: List[torch.Tensor]) -> None:
        """The kernels write parameters through raw pointers; tell autograd (and value caches) they changed."""
        fused.bump_param_epoch()
        try:
            torch.autograd.graph.increment_version(params)
        except Exception:  # pragma: no cover - older torch: per-tensor signature
            for p in params:
                try:
                    torch.autograd.graph.increment_version(p)
                except Exception:
                    return

    def _lr_tensor(self, gi: int, group: Dict[str, Any], device: torch.device) -> Optional[int]:
        """Device copy of the group's learning rate, refreshed only when the host value changed
        (outside CUDA-graph capture).  The kernels read the LR from it, so a captured step keeps
        following an LR schedule."""
        cache = self.__dict__.setdefault("_lr_cache", {})
        ent = cache.get(gi)
        lr = float(group["lr"])
        if ent is None:
            if torch.cuda.is_current_stream_capturing():
                return None
            ent = cache[gi] = [torch.full((), lr, dtype=torch.float32, device=device), lr]
        elif ent[1] != lr and not torch.cuda.is_current_stream_capturing():
            ent[0].fill_(lr)
            ent[1] = lr
        return ent[0].data_ptr()

    def sync_lr(self) -> None:
        """Push host-side LR changes (schedulers) to the device scalars; call before a graph replay."""
        for gi, g in enumerate(self.param_groups):
            if g["params"]:
                self._lr_tensor(gi, g, g["params"][0].device)

    def _consume_grad_mult(self) -> Optional[torch.Tensor]:
        if getattr(self, "_use_grad_mult", False):
            self._use_grad_mult = False
            return self._grad_mult
        return None


class FusedAdamW(_FusedBase):
    _adamw = 1
    writes_shadows = True        # its update kernel also emits the 16-bit copy of each parameter (desc.p_lp)

    def __init__(self, params, lr: float = 1e-3, betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, amsgrad: bool = False, maximize: bool = False, **_ignored: Any):
        if amsgrad:
            raise NotImplementedError("amsgrad is not implemented in the fused kernel")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, maximize=maximize))

    def _step_tensor(self, gi: int, params: List[torch.Tensor]) -> torch.Tensor:
        """One device-resident step counter per param group, shared by its parameters."""
        steps = self.__dict__.setdefault("_step_tensors", {})
        t = steps.get(gi)
        if t is None:
            seasoned = [p for p in params if len(self.state[p]) > 0]
            if seasoned:  # state came from a checkpoint
                t = torch.as_tensor(self.state[seasoned[0]]["step"], dtype=torch.float32).to(
                    params[0].device).reshape(()).clone()
                for p in seasoned:
                    self.state[p]["step"] = t
            else:
                t = torch.zeros((), dtype=torch.float32, device=params[0].device)
            steps[gi] = t
        return t

    def _init_group_state(self, gi: int, params: List[torch.Tensor]) -> None:
        step = self._step_tensor(gi, params)
        new = [p for p in params if len(self.state[p]) == 0]
        if not new:
            return
        _, ms = _flat_like(new)
        _, vs = _flat_like(new)
        for p, m, v in zip(new, ms, vs):
            self.state[p] = {"step": step, "exp_avg": m, "exp_avg_sq": v}

    @torch.no_grad()
    def step(self, closure: Optional[Callable] = None, found_inf: Optional[torch.Tensor] = None,
             grad_mult: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if grad_mult is None:
            grad_mult = self._consume_grad_mult()
        for gi, group in enumerate(self.param_groups):
            params = self._params_with_grad(group)
            if not params:
                continue
            self._init_group_state(gi, params)
            step_t = self._step_tensor(gi, params)
            grads = [p.grad for p in params]
            key = tuple(g.data_ptr() for g in grads) + tuple(p.data_ptr() for p in params)
            tl = self._tables.get(gi)
            if tl is None or tl.key != key:
                tl = self._tables[gi] = fused.TensorList(
                    params, grads, [self.state[p]["exp_avg"] for p in params],
                    [self.state[p]["exp_avg_sq"] for p in params], self._shadows(params))
            b1, b2 = group["betas"]
            hp = fused.AdamHParams(float(group["lr"]), float(b1), float(b2), float(group["eps"]),
                                   float(group["weight_decay"]), self._adamw, int(bool(group.get("maximize", False))),
                                   self._lr_tensor(gi, group, params[0].device))
            fused.step_advance(step_t, found_inf)
            fused.multi_adam(tl, grads[0].dtype, self.lp_dtype or torch.float16, hp, step_t, grad_mult, found_inf)
            self._bump_versions(params)
            fused.stamp_lp(params)
        return loss

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        super().load_state_dict(state_dict)
        self._tables.clear()
        self.__dict__["_step_tensors"] = {}
        for st in self.state.values():
            if "exp_avg" in st:
                st["exp_avg"] = st["exp_avg"].float()
                st["exp_avg_sq"] = st["exp_avg_sq"].float()
        for g in self.param_groups:
            for p in g["params"]:
                st = self.state.get(p)
                if st and "exp_avg" in st:  # state must share the parameter's memory layout
                    for k in ("exp_avg", "exp_avg_sq"):
                        if not fused.same_layout(st[k], p) or not fused.is_dense(st[k]):
                            st[k] = torch.empty_like(p, dtype=torch.float32).copy_(st[k])


class FusedAdam(FusedAdamW):
    """Adam with L2 regularisation (weight decay added to the gradient)."""
    _adamw = 0

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                 **kw: Any):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, **kw)


class FusedSGD(_FusedBase):
    def __init__(self, params, lr: float = 1e-3, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False, maximize: bool = False, **_ignored: Any):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                      nesterov=nesterov, maximize=maximize))

    @torch.no_grad()
    def step(self, closure: Optional[Callable] = None, found_inf: Optional[torch.Tensor] = None,
             grad_mult: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if grad_mult is None:
            grad_mult = self._consume_grad_mult()
        for gi, group in enumerate(self.param_groups):
            params = self._params_with_grad(group)
            if not params:
                continue
            mom = float(group["momentum"])
            # first-step semantics are per launch: split new / seasoned parameters
            for first, ps in ((True, [p for p in params if mom != 0 and "momentum_buffer" not in self.state[p]]),
                              (False, [p for p in params if mom == 0 or "momentum_buffer" in self.state[p]])):
                if not ps:
                    continue
                if first:
                    _, bufs = _flat_like(ps)
                    for p, b in zip(ps, bufs):
                        self.state[p]["momentum_buffer"] = b
                grads = [p.grad for p in ps]
                key = tuple(g.data_ptr() for g in grads) + tuple(p.data_ptr() for p in ps)
                tkey = gi * 2 + int(first)
                tl = self._tables.get(tkey)
                if tl is None or tl.key != key:
                    tl = self._tables[tkey] = fused.TensorList(
                        ps, grads, [self.state[p].get("momentum_buffer") for p in ps] if mom != 0 else None)
                hp = fused.SGDHParams(float(group["lr"]), mom, float(group["dampening"]), float(group["weight_decay"]),
                                      int(bool(group["nesterov"])), int(bool(group.get("maximize", False))),
                                      self._lr_tensor(gi, group, ps[0].device))
                fused.multi_sgd(tl, grads[0].dtype, torch.float16, hp, first, grad_mult, found_inf)
                self._bump_versions(ps)
        return loss

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        super().load_state_dict(state_dict)
        self._tables.clear()
        for g in self.param_groups:
            for p in g["params"]:
                st = self.state.get(p)
                if st and st.get("momentum_buffer") is not None:
                    b = st["momentum_buffer"].float()
                    if not fused.same_layout(b, p) or not fused.is_dense(b):
                        b = torch.empty_like(p, dtype=torch.float32).copy_(b)
                    st["momentum_buffer"] = b


def fused_equivalent(cls: Type[Optimizer]) -> Type[Optimizer]:
    """Map a stock optimizer class to its fused sm_100a implementation (identity if none)."""
    if cls is torch.optim.AdamW:
        return FusedAdamW
    if cls is torch.optim.Adam:
        return FusedAdam
    if cls is torch.optim.SGD:
        return FusedSGD
    return cls

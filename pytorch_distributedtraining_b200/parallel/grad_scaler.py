"""Dynamic loss scaling for fp16, plain and shard-aware.

Parity target: ``AMPConfig(init_scale=2.**14)`` handed to the facade
(Stoke-DDP.py:182-184); scaler semantics follow ``torch.amp.GradScaler``
(scale / unscale_ / step-skips-on-inf / update with growth tracker; state_dict keys
``scale, growth_factor, backoff_factor, growth_interval, _growth_tracker``) and, for
sharded engines, ``ShardedGradScaler`` (the inf flag must be OR-reduced across ranks
because each rank only sees its shard's gradients).

B200-first: the scale, the growth tracker and the found-inf flag are device scalars
and never visit the host on the training path.  ``step`` does not ``.item()`` the
flag: with a fused optimizer it is handed to the update kernel as a device-side
skip predicate; only foreign optimizers pay one host read.  On CUDA the unscale +
non-finite check is the multi-tensor kernel of ``ops`` (or is already folded into
the fused collective, in which case ``unscale_`` only merges flags).
"""
from __future__ import annotations

from collections import defaultdict
from enum import Enum
from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist

from . import comm

__all__ = ["GradScaler", "ShardedGradScaler"]


class _OptState(Enum):
    READY = 0
    UNSCALED = 1
    STEPPED = 2


class GradScaler:
    def __init__(self, init_scale: float = 2.0 ** 16, growth_factor: float = 2.0, backoff_factor: float = 0.5,
                 growth_interval: int = 2000, enabled: bool = True, device: Optional[torch.device] = None):
        self._enabled = enabled
        self._init_scale = float(init_scale)
        self._growth_factor = float(growth_factor)
        self._backoff_factor = float(backoff_factor)
        self._growth_interval = int(growth_interval)
        self._device = device
        self._scale: Optional[torch.Tensor] = None
        self._growth_tracker: Optional[torch.Tensor] = None
        self._found_inf: Optional[torch.Tensor] = None
        self._per_optimizer: Dict[int, Dict[str, Any]] = defaultdict(lambda: {"stage": _OptState.READY})

    # ------------------------------------------------------------------ lazy state
    def _lazy_init(self, device: torch.device) -> None:
        if self._scale is None:
            self._device = device
            self._scale = torch.full((), self._init_scale, dtype=torch.float32, device=device)
            self._growth_tracker = torch.full((), int(getattr(self, "_pending_tracker", 0) or 0), dtype=torch.int32,
                                              device=device)      # a checkpoint loaded before the first step
            self._pending_tracker = 0
            self._found_inf = torch.zeros((), dtype=torch.float32, device=device)

    def is_enabled(self) -> bool:
        return self._enabled

    def get_scale(self) -> float:
        if not self._enabled:
            return 1.0
        return self._init_scale if self._scale is None else float(self._scale.item())

    @property
    def scale_tensor(self) -> Optional[torch.Tensor]:
        return self._scale

    @property
    def found_inf_tensor(self) -> Optional[torch.Tensor]:
        return self._found_inf

    # ------------------------------------------------------------------ API
    def scale(self, outputs: Any) -> Any:
        if not self._enabled:
            return outputs
        if isinstance(outputs, torch.Tensor):
            self._lazy_init(outputs.device)
            return outputs * self._scale.to(outputs.dtype)
        return type(outputs)(self.scale(o) for o in outputs)

    def _grads_of(self, optimizer: torch.optim.Optimizer) -> List[torch.Tensor]:
        out = []
        for g in optimizer.param_groups:
            for p in g["params"]:
                if p.grad is not None:
                    out.append(p.grad)
        return out

    @torch.no_grad()
    def _unscale_grads(self, grads: List[torch.Tensor], inv_scale: torch.Tensor, found_inf: torch.Tensor) -> None:
        if not grads:
            return
        if grads[0].is_cuda:
            from ..ops import fused  # the sm_100a multi-tensor kernel; fails loudly if the library is missing
            fused.unscale_check_(grads, inv_scale, found_inf)
            return
        by_kind: Dict[Any, List[torch.Tensor]] = defaultdict(list)
        for g in grads:
            by_kind[(g.device, g.dtype)].append(g)
        for (dev, _dt), gs in by_kind.items():
            torch._amp_foreach_non_finite_check_and_unscale_(gs, found_inf.to(dev), inv_scale.to(dev))

    def unscale_(self, optimizer: torch.optim.Optimizer, grads: Optional[List[torch.Tensor]] = None,
                 already_unscaled: bool = False, flag_is_global: bool = False) -> None:
        """Divide gradients by the scale and record non-finite values.

        ``already_unscaled``: a fused collective did the arithmetic and wrote the flag
        into ``found_inf_tensor`` itself; only bookkeeping remains.  ``flag_is_global``: that
        flag has already been combined across ranks (skip the sharded scaler's all-reduce).
        """
        if not self._enabled:
            return
        st = self._per_optimizer[id(optimizer)]
        if st["stage"] is _OptState.UNSCALED:
            raise RuntimeError("unscale_() has already been called on this optimizer since the last update().")
        if st["stage"] is _OptState.STEPPED:
            raise RuntimeError("unscale_() is being called after step().")
        grads = self._grads_of(optimizer) if grads is None else grads
        if self._scale is None:
            self._lazy_init(grads[0].device if grads else torch.device("cpu"))
        if not already_unscaled:
            inv = self._scale.double().reciprocal().float()
            self._unscale_grads(grads, inv, self._found_inf)
        if not flag_is_global:
            self._sync_found_inf()
        st["stage"] = _OptState.UNSCALED

    def _sync_found_inf(self) -> None:
        """Hook for the sharded subclass."""

    def unscale_and_clip_fused(self, optimizer: Any, max_norm: Optional[float]) -> Optional[torch.Tensor]:
        """Fused path for optimizers exposing ``fused_unscale_clip``: one read of the
        gradients yields the inf flag and the squared norm; the unscale and the clip
        coefficient are applied inside the update kernel, so gradients are never
        rewritten.  Returns the (unscaled) total norm when clipping."""
        st = self._per_optimizer[id(optimizer)]
        if st["stage"] is not _OptState.READY:
            raise RuntimeError("unscale has already happened for this optimizer since the last update().")
        grads = self._grads_of(optimizer)
        if self._scale is None:
            self._lazy_init(grads[0].device if grads else torch.device("cpu"))
        inv = self._scale.reciprocal()
        total = optimizer.fused_unscale_clip(inv, max_norm, self._found_inf)
        self._sync_found_inf()
        st["stage"] = _OptState.UNSCALED
        return total

    def step(self, optimizer: torch.optim.Optimizer, *args: Any, **kwargs: Any) -> Optional[float]:
        if not self._enabled:
            return optimizer.step(*args, **kwargs)
        st = self._per_optimizer[id(optimizer)]
        if st["stage"] is _OptState.STEPPED:
            raise RuntimeError("step() has already been called since the last update().")
        if st["stage"] is _OptState.READY:
            self.unscale_(optimizer)
        ret = None
        if getattr(optimizer, "supports_device_skip", False):
            # fused optimizers take the flag as a device-side predicate: no host sync
            ret = optimizer.step(*args, found_inf=self._found_inf, **kwargs)
        elif float(self._found_inf.item()) == 0.0:
            ret = optimizer.step(*args, **kwargs)
        st["stage"] = _OptState.STEPPED
        return ret

    @torch.no_grad()
    def update(self, new_scale: Optional[float] = None) -> None:
        if not self._enabled or self._scale is None:
            return
        if new_scale is not None:
            self._scale.fill_(float(new_scale))
            self._found_inf.zero_()          # the fused update kernel clears it on the other path
        else:
            if self._scale.is_cuda:
                from ..ops import fused
                fused.scaler_update_(self._scale, self._growth_tracker, self._found_inf, self._growth_factor,
                                     self._backoff_factor, self._growth_interval)
            else:
                torch._amp_update_scale_(self._scale, self._growth_tracker, self._found_inf, self._growth_factor,
                                         self._backoff_factor, self._growth_interval)
        if not self._scale.is_cuda:
            self._found_inf.zero_()
        self._per_optimizer.clear()

    # ------------------------------------------------------------------ state
    def state_dict(self) -> Dict[str, Any]:
        if not self._enabled:
            return {}
        return {
            "scale": self.get_scale(),
            "growth_factor": self._growth_factor,
            "backoff_factor": self._backoff_factor,
            "growth_interval": self._growth_interval,
            "_growth_tracker": 0 if self._growth_tracker is None else int(self._growth_tracker.item()),
        }

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        if not self._enabled or not sd:
            return
        self._init_scale = float(sd["scale"])
        self._growth_factor = float(sd["growth_factor"])
        self._backoff_factor = float(sd["backoff_factor"])
        self._growth_interval = int(sd["growth_interval"])
        if self._scale is not None:
            self._scale.fill_(self._init_scale)
            self._growth_tracker.fill_(int(sd["_growth_tracker"]))
        else:
            self._pending_tracker = int(sd["_growth_tracker"])


class ShardedGradScaler(GradScaler):
    """Scaler for OSS / ShardedDDP / FSDP: OR the inf flag across the group."""

    def __init__(self, *args: Any, process_group: Optional[Any] = None, **kwargs: Any):
        super().__init__(*args, **kwargs)
        self.group = process_group if process_group is not None else (
            dist.group.WORLD if comm.is_initialized() else None)

    def _sync_found_inf(self) -> None:
        if comm.get_world_size(self.group) > 1:
            dist.all_reduce(self._found_inf, op=dist.ReduceOp.MAX, group=self.group)

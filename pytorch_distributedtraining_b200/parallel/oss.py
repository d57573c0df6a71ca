"""Optimizer-state sharding (ZeRO-1): every rank owns the optimizer state of a
disjoint subset of the parameters, steps only those, and publishes the updated
values to the other ranks.

Parity target: ``OSS(params=model.parameters(), optim=AdamW, lr=..., betas=...,
eps=..., weight_decay=...)`` (Fairscale-DDP.py:78-86) and the Stoke flag
``fairscale_oss=True`` with ``FairscaleOSSConfig(broadcast_fp16=True)``
(Stoke-DDP.py:197-199,249).  ``param_groups`` stays globally visible so LR
schedulers built on ``stoke_model.optimizer`` work (Stoke-DDP.py:300-306).

What is different from fairscale by design:
  * Ownership is a greedy size-balanced partition over parameters sorted by numel
    (deterministic on every rank, no communication).
  * With a 16-bit wire format the owner keeps an fp32 master copy of its shard, so
    the master weights are never rounded; every replica (owner included) computes
    with the same rounded values, which keeps replicas bit-identical.
  * On CUDA, when the sm_100a kernel library is attached (``attach_flat_engine``),
    step + publish are done by one fused kernel over a flat, owner-ordered buffer
    (``ops.collectives``); the code here is the portable semantic definition and the
    gloo/NCCL baseline.
"""
from __future__ import annotations

import copy
from collections import OrderedDict
from itertools import chain
from typing import Any, Callable, Dict, Iterable, List, Optional, Tuple, Type, Union

import torch
import torch.distributed as dist
from torch.optim import SGD, Optimizer

from . import comm

__all__ = ["OSS"]


def _to_cpu(obj: Any) -> Any:
    if isinstance(obj, torch.Tensor):
        return obj.detach().cpu()
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj


def _to_device(obj: Any, device: torch.device) -> Any:
    if isinstance(obj, torch.Tensor):
        return obj.to(device)
    if isinstance(obj, dict):
        return {k: _to_device(v, device) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_device(v, device) for v in obj)
    return obj


class OSS(Optimizer):
    """Wrap an arbitrary ``torch.optim.Optimizer`` class and shard its state.

    Args:
        params: iterable of parameters or param-group dicts (global view).
        optim: optimizer *class* to instantiate on the local shard.
        group: process group (default WORLD).
        broadcast_fp16: publish updated parameters as fp16 on the wire.
        broadcast_buffer_size: kept for API parity (ignored; buckets are per owner).
        broadcast_dtype: explicit wire dtype (overrides ``broadcast_fp16``), e.g. bf16.
        **default: optimizer hyper-parameters (``lr``, ``betas``...).
    """

    optim: Optimizer
    in_super_constructor: bool

    def __init__(
        self,
        params: Iterable,
        optim: Type[Optimizer] = SGD,
        group: Optional[Any] = None,
        broadcast_fp16: bool = False,
        broadcast_buffer_size: int = -1,
        broadcast_dtype: Optional[torch.dtype] = None,
        **default: Any,
    ):
        self.in_super_constructor = True
        super().__init__(params, default)
        self.in_super_constructor = False

        self.group = group if group is not None else (dist.group.WORLD if comm.is_initialized() else None)
        self.world_size = comm.get_world_size(self.group)
        self.rank = comm.get_rank(self.group)
        self.backend = dist.get_backend(self.group) if comm.is_initialized() else "none"
        self._optim_cls = optim
        self._optim_defaults = default
        self.broadcast_dtype: Optional[torch.dtype] = (
            broadcast_dtype if broadcast_dtype is not None else (torch.float16 if broadcast_fp16 else None)
        )
        self.broadcast_fp16 = broadcast_fp16

        # caches, all derived deterministically from param_groups
        self._partition: List[List[Dict[str, Any]]] = []
        self._param_to_rank: Dict[torch.Tensor, int] = {}
        self._param_to_index: Dict[int, int] = {}
        self._index_to_param: Dict[int, torch.Tensor] = {}
        self._per_device_params: Dict[torch.device, List[List[torch.Tensor]]] = {}
        self._all_states: List[Dict[str, Any]] = []  # filled by consolidate_state_dict
        self._master: Dict[torch.Tensor, torch.Tensor] = {}  # fp32 masters when wire dtype is 16-bit
        self._flat_engine = None  # set by attach_flat_engine (CUDA fused path)

        self.partition_parameters()
        local_groups = self._partition[self.rank]
        # An optimizer cannot be built over an empty list; give idle ranks a dummy.
        if sum(len(g["params"]) for g in local_groups) == 0:
            self._dummy = torch.nn.Parameter(torch.zeros(1), requires_grad=False)
            local_groups = [dict(local_groups[0] if local_groups else {}, params=[self._dummy])]
        self.optim = optim(local_groups, **default)
        self._default_device = self._any_device()
        if hasattr(self.optim, "fused_unscale_clip"):
            # the fused optimizers compute norm + inf flag in ONE read of the (owned) gradients and apply unscale / clip
            # inside the update kernel; exposing the entry point here lets the GradScaler / facade reach it through OSS
            self.fused_unscale_clip = self._fused_unscale_clip

    # ------------------------------------------------------------------ partition
    def _any_device(self) -> torch.device:
        for g in self.param_groups:
            for p in g["params"]:
                return p.device
        return torch.device("cpu")

    def _clear_cache(self) -> None:
        self._partition = []
        self._param_to_rank.clear()
        self._param_to_index.clear()
        self._index_to_param.clear()
        self._per_device_params.clear()

    def partition_parameters(self) -> List[List[Dict[str, Any]]]:
        """Greedy balanced partition: per group, largest parameter first, to the
        rank with the smallest running element count.  Pure function of
        ``param_groups`` so every rank computes the same answer.
        """
        if self._partition:
            return self._partition
        self._partition = [[] for _ in range(self.world_size)]
        sizes = [0] * self.world_size
        for group in self.param_groups:
            per_rank: List[List[torch.Tensor]] = [[] for _ in range(self.world_size)]
            order = sorted(range(len(group["params"])), key=lambda i: (-group["params"][i].numel(), i))
            for i in order:
                p = group["params"][i]
                r = min(range(self.world_size), key=lambda k: (sizes[k], k))
                per_rank[r].append(p)
                # frozen params still get an owner (state_dict layout), but cost nothing
                sizes[r] += p.numel() if p.requires_grad else 0
            for r, plist in enumerate(per_rank):
                # keep the original order inside a rank's slice (stable state_dict mapping)
                pos = {id(p): k for k, p in enumerate(group["params"])}
                plist.sort(key=lambda p: pos[id(p)])
                g = {k: v for k, v in group.items() if k != "params"}
                g["params"] = plist
                self._partition[r].append(g)
        return self._partition

    @property
    def param_to_rank(self) -> Dict[torch.Tensor, int]:
        if not self._param_to_rank:
            for r, groups in enumerate(self.partition_parameters()):
                for g in groups:
                    for p in g["params"]:
                        self._param_to_rank[p] = r
        return self._param_to_rank

    @property
    def param_to_index(self) -> Dict[int, int]:
        """``id(param)`` -> global index in torch ``state_dict`` numbering."""
        if not self._param_to_index:
            for i, p in enumerate(chain.from_iterable(g["params"] for g in self.param_groups)):
                self._param_to_index[id(p)] = i
        return self._param_to_index

    @property
    def index_to_param(self) -> Dict[int, torch.Tensor]:
        if not self._index_to_param:
            for i, p in enumerate(chain.from_iterable(g["params"] for g in self.param_groups)):
                self._index_to_param[i] = p
        return self._index_to_param

    @property
    def per_device_params(self) -> Dict[torch.device, List[List[torch.Tensor]]]:
        """device -> rank -> owned parameters (used by ShardedDataParallel)."""
        if not self._per_device_params:
            for g in self.param_groups:
                for p in g["params"]:
                    dev = p.device
                    if dev not in self._per_device_params:
                        self._per_device_params[dev] = [[] for _ in range(self.world_size)]
                    self._per_device_params[dev][self.param_to_rank[p]].append(p)
        return self._per_device_params

    def local_params(self) -> List[torch.Tensor]:
        return [p for g in self.partition_parameters()[self.rank] for p in g["params"]]

    def refresh_trainable(self) -> None:
        """Call after flipping ``requires_grad``: re-balances nothing (ownership is
        sticky so optimizer state stays valid) but drops derived caches."""
        self._per_device_params.clear()

    # ------------------------------------------------------------------ step
    def _sync_param_groups(self, source: List[Dict], destination: List[Dict]) -> None:
        """Copy hyper-parameters (lr set by a scheduler, ...) between group lists."""
        for src, dst in zip(source, destination):
            for k in src.keys():
                if k != "params":
                    dst[k] = src[k]

    def step(self, closure: Optional[Callable[[], float]] = None, **kwargs: Any) -> Optional[float]:
        if self._flat_engine is not None:
            return self._flat_engine.step(closure=closure, **kwargs)
        self._sync_param_groups(self.param_groups, self.optim.param_groups)
        self._restore_masters()
        loss = self.optim.step(closure=closure, **kwargs) if closure is not None else self.optim.step(**kwargs)
        self._stash_masters()
        self._broadcast_params()
        self._sync_param_groups(self.optim.param_groups, self.param_groups)
        return loss

    def zero_grad(self, set_to_none: bool = True) -> None:
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is not None:
                    if set_to_none:
                        p.grad = None
                    else:
                        p.grad.detach_()
                        p.grad.zero_()

    # fp32 master handling for 16-bit wire formats -----------------------------
    def _wire_active(self, p: torch.Tensor) -> bool:
        return self.broadcast_dtype is not None and p.dtype == torch.float32 and self.world_size > 1

    @torch.no_grad()
    def _restore_masters(self) -> None:
        for p, m in self._master.items():
            p.copy_(m)

    @torch.no_grad()
    def _stash_masters(self) -> None:
        if self.broadcast_dtype is None or self.world_size == 1:
            return
        for p in self.local_params():
            if self._wire_active(p):
                m = self._master.get(p)
                if m is None:
                    self._master[p] = p.detach().clone()
                else:
                    m.copy_(p)

    @torch.no_grad()
    def _broadcast_params(self) -> None:
        """Owner -> everyone, one flat bucket per (owner, device, dtype) (K3)."""
        if self.world_size == 1:
            return
        handles: List[Tuple[Any, torch.Tensor, List[torch.Tensor]]] = []
        for _dev, per_rank in self.per_device_params.items():
            for owner, plist in enumerate(per_rank):
                plist = [p for p in plist if p.requires_grad or True]
                if not plist:
                    continue
                by_dtype: Dict[torch.dtype, List[torch.Tensor]] = OrderedDict()
                for p in plist:
                    by_dtype.setdefault(p.dtype, []).append(p)
                for dtype, ps in by_dtype.items():
                    wire = self.broadcast_dtype if (self.broadcast_dtype is not None and dtype == torch.float32) else dtype
                    flat = torch.cat([p.detach().reshape(-1) for p in ps]).to(wire)
                    h = dist.broadcast(flat, src=comm.get_global_rank(self.group, owner), group=self.group, async_op=True)
                    handles.append((h, flat, ps))
        for h, flat, ps in handles:
            h.wait()
            off = 0
            for p in ps:
                n = p.numel()
                p.copy_(flat[off:off + n].view_as(p))
                off += n

    # ------------------------------------------------------------------ clipping
    @torch.no_grad()
    def clip_grad_norm(
        self,
        max_norm: Union[float, int],
        norm_type: Union[float, int] = 2.0,
        filter_params_fn: Optional[Callable[[Any], Any]] = None,
    ) -> torch.Tensor:
        """Shard-aware global-norm clipping (Stoke-DDP.py:253 via the facade).

        Each rank reduces over the parameters it owns (with gradient sharding those
        are the only ones whose reduced gradient it holds), the partial ``norm**p``
        is summed across ranks (MAX for the inf-norm), and every gradient present on
        this rank is scaled.
        """
        if self._flat_engine is not None:
            return self._flat_engine.clip_grad_norm(max_norm, norm_type)
        max_norm = float(max_norm)
        norm_type = float(norm_type)
        if norm_type == 2.0 and filter_params_fn is None and self._fused_clip_usable():
            return self._fused_unscale_clip(None, max_norm, None)
        local = [p for p in self.local_params() if p.grad is not None]
        if filter_params_fn is not None:
            local = list(filter_params_fn(local))
        device = self._default_device
        if norm_type == float("inf"):
            total = torch.zeros((), device=device)
            for p in local:
                total = torch.maximum(total, p.grad.detach().abs().max().to(device))
            if self.world_size > 1:
                dist.all_reduce(total, op=dist.ReduceOp.MAX, group=self.group)
        else:
            total = torch.zeros((), device=device, dtype=torch.float32)
            for p in local:
                total += torch.linalg.vector_norm(p.grad.detach().float(), norm_type).to(device) ** norm_type
            if self.world_size > 1:
                dist.all_reduce(total, op=dist.ReduceOp.SUM, group=self.group)
            total = total ** (1.0 / norm_type)
        clip_coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is not None:
                    p.grad.detach().mul_(clip_coef.to(p.grad.device, p.grad.dtype))
        return total

    def _fused_clip_usable(self) -> bool:
        from ..ops import dispatch
        if "fused_unscale_clip" not in self.__dict__ or not dispatch.enabled("oss_fused_clip"):
            return False
        if self.world_size > 1 and not dispatch.enabled("oss_fused_clip_multirank"):
            return False
        grads = [p.grad for p in self.local_params() if p.grad is not None]
        return bool(grads) and all(g.is_cuda for g in grads)

    def _fused_unscale_clip(self, inv_scale: Optional[torch.Tensor], max_norm: Optional[float],
                            found_inf: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        """Shard-aware norm / inf check with ONE read of the gradients this rank owns (two launches instead of four per
        parameter): the partial sum of squares and the flag are summed across ranks, the resulting multiplier (1/scale
        and the clip coefficient) is applied inside the inner optimizer's update kernel — gradients are not rewritten.
        Returns the global gradient norm (device scalar) when clipping."""
        reduce = None
        if self.world_size > 1:
            def reduce(buf: torch.Tensor) -> None:      # [sum g^2, non-finite flag]
                dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
                buf[1:2].clamp_(max=1.0)
        return self.optim.fused_unscale_clip(inv_scale, max_norm, found_inf, extra_sumsq_reduce=reduce)

    # ------------------------------------------------------------------ state dict
    @property
    def _local_index_map(self) -> List[int]:
        """local optimizer param index -> global index."""
        out = []
        for g in self.partition_parameters()[self.rank]:
            for p in g["params"]:
                out.append(self.param_to_index[id(p)])
        return out

    def local_state_dict(self) -> Dict[str, Any]:
        """This rank's shard, in the inner optimizer's own numbering, plus the map
        to global indices."""
        self._restore_masters_into_state_view()
        sd = self.optim.state_dict()
        return {"state": sd["state"], "param_groups": sd["param_groups"], "global_index": self._local_index_map}

    def _restore_masters_into_state_view(self) -> None:
        # masters live outside the inner optimizer; nothing to fold in (they are
        # parameters, not optimizer state) — kept as a hook for the flat engine.
        return

    def consolidate_state_dict(self, recipient_rank: int = 0) -> None:
        """Gather every rank's shard of the optimizer state onto ``recipient_rank``
        (``-1``: onto all ranks).  Collective: every rank must call it (K12)."""
        if self._flat_engine is not None:
            local = self._flat_engine.local_state_dict()
        else:
            local = self.local_state_dict()
        local = _to_cpu(local)
        if self.world_size == 1:
            self._all_states = [local]
            return
        if recipient_rank == -1:
            self._all_states = comm.all_gather_object(local, group=self.group)
            return
        if self.backend == "nccl":
            # gather_object on NCCL stages through the current CUDA device; all_gather keeps it simple
            gathered = comm.all_gather_object(local, group=self.group)
            self._all_states = gathered if self.rank == recipient_rank else []
        else:
            out = [None] * self.world_size if self.rank == recipient_rank else None
            dist.gather_object(local, out, dst=comm.get_global_rank(self.group, recipient_rank), group=self.group)
            self._all_states = out if self.rank == recipient_rank else []

    def state_dict(self) -> Dict[str, Any]:
        """Full optimizer state in the *plain torch optimizer layout* (state keyed by
        global parameter index; ``param_groups[i]['params']`` lists global indices),
        so it loads into an un-sharded ``torch.optim`` instance of the same class and
        into an OSS of any world size.  Requires ``consolidate_state_dict`` first
        when world_size > 1.
        """
        if self.world_size == 1 and not self._all_states:
            self.consolidate_state_dict()
        if not self._all_states:
            raise RuntimeError(
                "Optimizer state has not been consolidated on this rank. "
                "Call consolidate_state_dict(recipient_rank) on every rank first."
            )
        state: Dict[int, Any] = {}
        for shard in self._all_states:
            gmap = shard["global_index"]
            for local_idx, s in shard["state"].items():
                if int(local_idx) < len(gmap):
                    state[gmap[int(local_idx)]] = s
        groups = []
        start = 0
        for g in self.param_groups:
            gd = {k: v for k, v in g.items() if k != "params"}
            gd["params"] = list(range(start, start + len(g["params"])))
            start += len(g["params"])
            groups.append(gd)
        return {"state": dict(sorted(state.items())), "param_groups": groups}

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        """Load a *full* torch-layout state dict; keep only the local shard."""
        state_dict = copy.copy(state_dict)
        for g_new, g_saved in zip(self.param_groups, state_dict["param_groups"]):
            for k, v in g_saved.items():
                if k != "params":
                    g_new[k] = v
        self._sync_param_groups(self.param_groups, self.optim.param_groups)
        if self._flat_engine is not None:
            self._flat_engine.load_full_state(state_dict["state"])
            return
        gmap = self._local_index_map
        local_state = {}
        for local_idx, gidx in enumerate(gmap):
            s = state_dict["state"].get(gidx, state_dict["state"].get(str(gidx)))
            if s is not None:
                p = self.index_to_param[gidx]
                local_state[local_idx] = {
                    k: (v.to(p.device) if isinstance(v, torch.Tensor) and v.dim() > 0 else copy.deepcopy(v))
                    for k, v in s.items()
                }
        inner = self.optim.state_dict()
        inner["state"] = local_state
        self.optim.load_state_dict(inner)
        self._master.clear()
        self._all_states = []

    # torch.optim.Optimizer plumbing ---------------------------------------------
    def add_param_group(self, param_group: dict) -> None:
        super().add_param_group(param_group)
        if not self.in_super_constructor:
            self._clear_cache()
            self.partition_parameters()
            new_local = self._partition[self.rank][-1]
            if new_local["params"]:
                self.optim.add_param_group(new_local)

    @property
    def supports_device_skip(self) -> bool:
        """True when the inner optimizer takes the GradScaler's inf flag as a device predicate."""
        return bool(getattr(self.optim, "supports_device_skip", False))

    @property
    def inner_state(self):
        return self.optim.state

    def attach_flat_engine(self, engine: Any) -> None:
        """Route step/clip/state through the fused flat-buffer CUDA engine."""
        self._flat_engine = engine

    def __repr__(self) -> str:
        return (f"OSS(optim={self._optim_cls.__name__}, world_size={self.world_size}, rank={self.rank}, "
                f"wire={self.broadcast_dtype}, owned={sum(p.numel() for p in self.local_params())})")

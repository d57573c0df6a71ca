"""Full parameter sharding (ZeRO-3).

Parity target: the engine behind Stoke's ``fairscale_fsdp`` flag and
``FairscaleFSDPConfig`` (BASELINE.json config 4, ViT-B/16 bf16).
Neither reference script enables it, but it is on the facade surface they use
(Stoke-DDP.py:240-254).

Design (functional, not storage-swizzling of ``nn.Parameter``):
  * every wrapped unit flattens its own (non-nested) parameters into one padded flat
    tensor and keeps ``chunk(world)[rank]`` as a single fp32 ``nn.Parameter``
    (``flat_param``) — that is the only thing the optimizer ever sees;
  * forward runs ``full = Unshard(flat_param)`` — an autograd function whose forward
    is the **all-gather** (in ``compute_dtype``) and whose backward is the
    **reduce-scatter** of the full gradient (pre-divide → reduce → post-divide →
    fp32 shard gradient).  The unit's modules see ``full.split(...)`` views, so the
    per-parameter gradients are assembled by a single ``cat`` in autograd;
  * ``reshard_after_forward`` frees the gathered storage after forward and re-gathers
    it in a pre-backward hook, writing into the same storage so tensors saved by
    autograd stay valid;
  * ``no_sync()`` accumulates the unsharded gradient locally and reduces once.

On CUDA with the kernel library the all-gather / reduce-scatter are the
peer-memory kernels (``parallel.symm.FSDPPeerEngine``): the gather casts fp32 shards to the
compute dtype on the fly, writes into a small ring of gathered-parameter slots and is
PREFETCHED one unit ahead on its own stream; the reduce-scatter is fused with the scale and
the fp32 cast (W5), reads the gradient pieces from one symmetric staging buffer and runs on a
third stream under the backward of the units below.  The gather-inside-the-GEMM kernel (W4,
``ops.gemm.gemm_gather``) is NOT used here: with the gather prefetched a full unit ahead it
measured no faster than gather + GEMM.  With gloo/NCCL the ``dist``
collectives are used — that is the baseline and the CPU test path.
"""
from __future__ import annotations

import contextlib
from typing import Any, Callable, Dict, Generator, Iterator, List, Optional, Set, Tuple

import torch
import torch.distributed as dist
import torch.nn as nn

from . import comm

__all__ = ["FullyShardedDataParallel", "auto_wrap", "default_auto_wrap_policy"]


def _all_gather_into(out: torch.Tensor, shard: torch.Tensor, group: Any, world: int) -> None:
    if world == 1:
        out.copy_(shard)
        return
    backend = dist.get_backend(group)
    if backend == "nccl":
        dist.all_gather_into_tensor(out, shard, group=group)
    else:
        dist.all_gather(list(out.chunk(world)), shard, group=group)


def _reduce_scatter_into(out: torch.Tensor, full: torch.Tensor, group: Any, world: int) -> None:
    if world == 1:
        out.copy_(full)
        return
    backend = dist.get_backend(group)
    if backend == "nccl":
        dist.reduce_scatter_tensor(out, full, group=group)
    else:
        # gloo has no reduce_scatter: all-reduce then slice (plumbing path only)
        dist.all_reduce(full, group=group)
        out.copy_(full.chunk(world)[dist.get_rank(group)])


class _Unshard(torch.autograd.Function):
    """forward: shard -> per-parameter views of the gathered flat tensor; backward: the per-parameter gradients ->
    shard gradient.  Returning the VIEWS (instead of the flat tensor, split afterwards) hands the backward the
    individual gradient pieces: the peer-memory path copies them straight into its symmetric staging buffer and
    reduce-scatters from there — one copy in total, where ``cat`` + stage would be two."""

    @staticmethod
    def forward(ctx, flat_shard: torch.Tensor, unit: "FullyShardedDataParallel"):  # type: ignore[override]
        ctx.unit = unit
        ctx.set_materialize_grads(False)
        # a fresh alias each call: clean autograd history, same storage (so the
        # re-gather before backward refills what autograd saved)
        return unit._split(unit._gather_full().detach())

    @staticmethod
    def backward(ctx, *grads: Optional[torch.Tensor]):  # type: ignore[override]
        unit: FullyShardedDataParallel = ctx.unit
        return unit._reduce_grad(grads), None


class FullyShardedDataParallel(nn.Module):
    def __init__(
        self,
        module: nn.Module,
        process_group: Optional[Any] = None,
        reshard_after_forward: bool = True,
        mixed_precision: bool = False,
        fp32_reduce_scatter: bool = False,
        flatten_parameters: bool = True,
        move_params_to_cpu: bool = False,
        compute_dtype: Optional[torch.dtype] = None,
        buffer_dtype: Optional[torch.dtype] = None,
        move_grads_to_cpu: Optional[bool] = None,
        bucket_cap_mb: int = 25,
        compute_device: Optional[torch.device] = None,
        no_broadcast_optim_state: Optional[bool] = False,
        clear_autocast_cache: bool = False,
        force_input_to_fp32: bool = False,
        verbose: bool = False,
        gradient_predivide_factor: Optional[float] = None,
        gradient_postdivide_factor: Optional[float] = None,
        sync_models_at_startup: bool = True,
    ):
        super().__init__()
        self.process_group = process_group if process_group is not None else (
            dist.group.WORLD if comm.is_initialized() else None)
        self.world_size = comm.get_world_size(self.process_group)
        self.rank = comm.get_rank(self.process_group)
        self.reshard_after_forward = reshard_after_forward and self.world_size > 1
        self.mixed_precision = mixed_precision
        self.fp32_reduce_scatter = fp32_reduce_scatter
        self.compute_dtype = compute_dtype or (torch.float16 if mixed_precision else torch.float32)
        self.buffer_dtype = buffer_dtype
        self.force_input_to_fp32 = force_input_to_fp32
        self.verbose = verbose
        if move_params_to_cpu or move_grads_to_cpu:
            raise NotImplementedError("CPU offload is not implemented: a B200 holds 180 GB; shard instead")
        self.gradient_predivide_factor = float(
            gradient_predivide_factor if gradient_predivide_factor is not None
            else self._default_predivide(self.world_size))
        self.gradient_postdivide_factor = float(
            gradient_postdivide_factor if gradient_postdivide_factor is not None
            else self.world_size / self.gradient_predivide_factor)
        self._require_backward_grad_sync = True
        self._peer_engine: Any = None

        self.module = module
        if sync_models_at_startup and self.world_size > 1:
            tensors = [p.data for p in self._own_params(module)] + [b.data for b in module.buffers(recurse=True)]
            comm.broadcast_coalesced(tensors, src=0, group=self.process_group)
        self._flatten()
        self._full: Optional[torch.Tensor] = None       # gathered flat tensor (compute dtype)
        self._full_valid = False
        self._grad_accum_full: Optional[torch.Tensor] = None  # unsharded grads under no_sync
        self._in_backward = False
        self._is_root: Optional[bool] = None
        self._summoned = False

    @staticmethod
    def _default_predivide(world: int) -> float:
        f = 1
        while world % f == 0 and world / f > f:
            f *= 2
        return float(f)

    # ------------------------------------------------------------------ flatten
    @staticmethod
    def _own_params(module: nn.Module) -> List[nn.Parameter]:
        """Parameters of ``module`` that are not inside a nested FSDP unit."""
        out: List[nn.Parameter] = []
        seen: Set[int] = set()

        def visit(m: nn.Module) -> None:
            for p in m._parameters.values():
                if p is not None and id(p) not in seen:
                    seen.add(id(p))
                    out.append(p)
            for c in m.children():
                if not isinstance(c, FullyShardedDataParallel):
                    visit(c)

        visit(module)
        return out

    def _flatten(self) -> None:
        infos: List[Tuple[nn.Module, str, torch.Size, int, str]] = []
        seen: Dict[int, int] = {}
        tensors: List[torch.Tensor] = []
        shared: List[Tuple[nn.Module, str, int]] = []

        def visit(m: nn.Module, prefix: str) -> None:
            for name, p in list(m._parameters.items()):
                if p is None:
                    continue
                if id(p) in seen:
                    shared.append((m, name, seen[id(p)]))
                    continue
                seen[id(p)] = len(infos)
                infos.append((m, name, p.shape, p.numel(), prefix + name))
                tensors.append(p.detach().reshape(-1).float())
            for cname, c in m.named_children():
                if not isinstance(c, FullyShardedDataParallel):
                    visit(c, prefix + cname + ".")

        visit(self.module, "")
        self._infos = infos
        self._shared = shared
        self._numels = [i[3] for i in infos]
        self._total = sum(self._numels)
        requires_grad = any(getattr(m, "_parameters")[n].requires_grad for m, n, *_ in infos) if infos else False
        device = tensors[0].device if tensors else torch.device("cpu")
        # pad so each rank's shard is a multiple of 128 elements (16 B vectors, whole kernel tiles)
        quantum = 128 * self.world_size
        self._padded = max((self._total + quantum - 1) // quantum * quantum, quantum) if infos else 0
        self._shard_numel = self._padded // self.world_size if infos else 0
        if infos:
            flat = torch.zeros(self._padded, dtype=torch.float32, device=device)
            torch.cat(tensors, out=flat[: self._total])
            shard = flat[self.rank * self._shard_numel: (self.rank + 1) * self._shard_numel].clone()
            self.flat_param = nn.Parameter(shard, requires_grad=requires_grad)
            # replace the originals by plain attributes (filled with views at forward time)
            for m, name, *_ in infos:
                del m._parameters[name]
                setattr(m, name, None)
            for m, name, _ in shared:
                del m._parameters[name]
                setattr(m, name, None)
        else:
            self.flat_param = None  # type: ignore[assignment]

    # ------------------------------------------------------------------ gather / reshard
    def attach_peer_engine(self, engine: Any) -> None:
        self._peer_engine = engine

    def _alloc_full(self) -> torch.Tensor:
        if self._full is None:
            self._full = torch.empty(self._padded, dtype=self.compute_dtype, device=self.flat_param.device)
        elif self._full.untyped_storage().size() == 0:
            self._full.untyped_storage().resize_(self._padded * self._full.element_size())
        return self._full

    @torch.no_grad()
    def _gather_full(self) -> torch.Tensor:
        full = self._alloc_full()
        if self._full_valid:
            return full
        shard = self.flat_param.detach()
        with torch.autograd._unsafe_preserve_version_counter(full):
            if self._peer_engine is not None:
                # peer kernel: reads the fp32 shard, casts to the compute dtype on the fly (W4 producer half)
                self._peer_engine.all_gather(full, shard)
            else:
                if shard.dtype != self.compute_dtype:
                    shard = shard.to(self.compute_dtype)
                _all_gather_into(full, shard, self.process_group, self.world_size)
        self._full_valid = True
        return full

    def _free_full(self) -> None:
        if self._full is not None and not getattr(self, "_full_persistent", False):
            self._full.untyped_storage().resize_(0)
        self._full_valid = False

    def _split(self, full: torch.Tensor) -> Tuple[torch.Tensor, ...]:
        pad = self._padded - self._total
        return tuple(full.split(self._numels + ([pad] if pad else [])))[: len(self._numels)]

    def _install_views(self, full) -> None:
        """``full``: the gathered flat tensor, or the tuple of per-parameter views of it."""
        views = full if isinstance(full, tuple) else self._split(full)
        for (m, name, shape, _n, _fq), v in zip(self._infos, views):
            setattr(m, name, v.view(shape))
        for m, name, idx in self._shared:
            src_m, src_name = self._infos[idx][0], self._infos[idx][1]
            setattr(m, name, getattr(src_m, src_name))

    def _cat_grads(self, grads) -> torch.Tensor:
        """Per-parameter gradient pieces (``None``: no gradient) -> flat padded gradient in the compute dtype."""
        ref = next((g for g in grads if g is not None), None)
        dt = ref.dtype if ref is not None else self.compute_dtype
        dev = ref.device if ref is not None else self.flat_param.device
        pieces = [(g.reshape(-1) if g is not None else torch.zeros(n, dtype=dt, device=dev)) for g, n in zip(grads, self._numels)]
        pad = self._padded - self._total
        if pad:
            pieces.append(torch.zeros(pad, dtype=dt, device=dev))
        return torch.cat(pieces)

    @torch.no_grad()
    def _reduce_grad(self, grads) -> Optional[torch.Tensor]:
        """Post-backward for this unit: (accumulate) → pre-divide → reduce-scatter →
        post-divide → fp32 shard gradient.  ``grads``: the per-parameter gradient pieces."""
        if (self._peer_engine is not None and self._require_backward_grad_sync and self._grad_accum_full is None
                and not self.fp32_reduce_scatter and hasattr(self._peer_engine, "reduce_scatter_pieces")):
            # peer path: pieces -> symmetric staging (the only copy) -> reduce-scatter + scale + fp32 cast on the
            # reduce stream, overlapped with the backward of the units below; the shard gradient is attached to
            # flat_param.grad at the end of backward (nothing flows through autograd here)
            if self._full is not None:
                self._install_views(self._full.detach())
            if self.reshard_after_forward:
                self._free_full()
            self._peer_engine.reduce_scatter_pieces(grads, 1.0 / self.world_size)
            return None
        grad_full = self._cat_grads(grads)
        # the module attributes still point at views of this iteration's autograd graph; swap them
        # for history-free views so the graph (and its AccumulateGrad nodes, which would otherwise be
        # re-used by the next iteration and break CUDA-graph capture) can be freed
        if self._full is not None:
            self._install_views(self._full.detach())
        if self.reshard_after_forward:
            self._free_full()
        if not self._require_backward_grad_sync:
            if self._grad_accum_full is None:
                self._grad_accum_full = grad_full.detach().clone()
            else:
                self._grad_accum_full.add_(grad_full)
            return None
        if self._grad_accum_full is not None:
            grad_full = grad_full + self._grad_accum_full
            self._grad_accum_full = None
        if self.fp32_reduce_scatter and grad_full.dtype != torch.float32:
            grad_full = grad_full.float()
        if self._peer_engine is not None:
            # peer kernel: sum across ranks, scale by 1/world and write the fp32 shard gradient (W5)
            out = torch.empty(self._shard_numel, dtype=torch.float32, device=grad_full.device)
            self._peer_engine.reduce_scatter(out, grad_full, 1.0 / self.world_size)
            return out.to(self.flat_param.dtype)
        out = torch.empty(self._shard_numel, dtype=grad_full.dtype, device=grad_full.device)
        if self.gradient_predivide_factor > 1:
            grad_full = grad_full / self.gradient_predivide_factor
        elif not grad_full.is_contiguous():
            grad_full = grad_full.contiguous()
        _reduce_scatter_into(out, grad_full, self.process_group, self.world_size)
        if self.gradient_postdivide_factor > 1:
            out.div_(self.gradient_postdivide_factor)
        return out.to(self.flat_param.dtype)

    # ------------------------------------------------------------------ forward
    def _cast_inputs(self, args: Tuple, kwargs: Dict) -> Tuple[Tuple, Dict]:
        if not self.mixed_precision and not self.force_input_to_fp32:
            return args, kwargs
        tgt = torch.float32 if self.force_input_to_fp32 else self.compute_dtype

        def cast(x: Any) -> Any:
            if isinstance(x, torch.Tensor) and x.is_floating_point():
                return x.to(tgt)
            return x

        return tuple(cast(a) for a in args), {k: cast(v) for k, v in kwargs.items()}

    def forward(self, *args: Any, **kwargs: Any) -> Any:
        if self._is_root is None:
            self._set_root()
        if self.flat_param is None:
            return self.module(*args, **kwargs)
        if self._is_root:
            args, kwargs = self._cast_inputs(args, kwargs)
        if not self._summoned:
            self._full_valid = False  # the optimizer may have moved the shard since the last gather
        if torch.is_grad_enabled() and self.flat_param.requires_grad:
            full = _Unshard.apply(self.flat_param, self)
        else:
            full = self._gather_full()
        self._install_views(full)
        out = self.module(*args, **kwargs)
        if self.reshard_after_forward and not self._summoned:
            needs_bwd = torch.is_grad_enabled() and self.flat_param.requires_grad
            self._free_full()
            if needs_bwd:
                out = self._register_pre_backward(out)
        return out

    def _register_pre_backward(self, out: Any) -> Any:
        """Re-gather just before autograd reaches this unit's outputs."""
        fired = [False]

        def hook(_grad: torch.Tensor) -> None:
            if not fired[0]:
                fired[0] = True
                self._bwd_gather = True          # tells a prefetching engine which way the pass is moving
                try:
                    self._gather_full()
                finally:
                    self._bwd_gather = False

        def visit(x: Any) -> Any:
            if isinstance(x, torch.Tensor) and x.requires_grad:
                x.register_hook(hook)
            elif isinstance(x, (list, tuple)):
                for y in x:
                    visit(y)
            elif isinstance(x, dict):
                for y in x.values():
                    visit(y)
            return x

        return visit(out)

    def _set_root(self) -> None:
        self._is_root = True
        for m in self.module.modules():
            if isinstance(m, FullyShardedDataParallel) and m is not self:
                m._is_root = False

    # ------------------------------------------------------------------ API
    def _fsdp_units(self) -> List["FullyShardedDataParallel"]:
        return [m for m in self.modules() if isinstance(m, FullyShardedDataParallel)]

    @contextlib.contextmanager
    def no_sync(self) -> Generator:
        units = self._fsdp_units()
        old = [u._require_backward_grad_sync for u in units]
        for u in units:
            u._require_backward_grad_sync = False
        try:
            yield
        finally:
            for u, o in zip(units, old):
                u._require_backward_grad_sync = o

    @contextlib.contextmanager
    def summon_full_params(self, recurse: bool = True, writeback: bool = True) -> Generator:
        """Materialise full fp32 parameters as real attributes; optionally write local
        modifications back into the shard on exit."""
        units = self._fsdp_units() if recurse else [self]
        fulls = []
        for u in units:
            if u.flat_param is None:
                fulls.append(None)
                continue
            full = torch.empty(u._padded, dtype=torch.float32, device=u.flat_param.device)
            with torch.no_grad():
                _all_gather_into(full, u.flat_param.detach(), u.process_group, u.world_size)
            u._install_views(full)
            u._summoned = True
            fulls.append(full)
        try:
            yield
        finally:
            for u, full in zip(units, fulls):
                if full is None:
                    continue
                if writeback:
                    with torch.no_grad():
                        u.flat_param.copy_(full[u.rank * u._shard_numel: (u.rank + 1) * u._shard_numel])
                u._summoned = False
                u._full_valid = False

    @torch.no_grad()
    def clip_grad_norm_(self, max_norm: float, norm_type: float = 2.0) -> torch.Tensor:
        """Global norm over the *sharded* gradients of every unit (all ranks call)."""
        grads = [u.flat_param.grad for u in self._fsdp_units() if u.flat_param is not None and u.flat_param.grad is not None]
        device = grads[0].device if grads else (self.flat_param.device if self.flat_param is not None else torch.device("cpu"))
        norm_type = float(norm_type)
        if norm_type == float("inf"):
            total = torch.zeros((), device=device)
            for g in grads:
                total = torch.maximum(total, g.abs().max())
            if self.world_size > 1:
                dist.all_reduce(total, op=dist.ReduceOp.MAX, group=self.process_group)
        else:
            total = torch.zeros((), device=device)
            for g in grads:
                total += torch.linalg.vector_norm(g.float(), norm_type) ** norm_type
            if self.world_size > 1:
                dist.all_reduce(total, group=self.process_group)
            total = total ** (1.0 / norm_type)
        coef = torch.clamp(float(max_norm) / (total + 1e-6), max=1.0)
        for g in grads:
            g.mul_(coef.to(g.dtype))
        return total

    # state dicts ----------------------------------------------------------------
    def _named_full_tensors(self, prefix: str = "") -> Iterator[Tuple[str, torch.Tensor]]:
        """(fully-qualified original name, full fp32 tensor) for this unit only."""
        if self.flat_param is None:
            return
        full = torch.empty(self._padded, dtype=torch.float32, device=self.flat_param.device)
        with torch.no_grad():
            _all_gather_into(full, self.flat_param.detach(), self.process_group, self.world_size)
        for (m, name, shape, n, fq), v in zip(self._infos, full[: self._total].split(self._numels)):
            yield prefix + fq, v.view(shape).clone()

    def state_dict(self, *args: Any, destination=None, prefix: str = "", keep_vars: bool = False):  # type: ignore[override]
        """Full, un-flattened, un-prefixed state dict identical in layout to the
        wrapped module's own (every rank must call: it gathers)."""
        out: Dict[str, torch.Tensor] = destination if destination is not None else {}
        self._collect_state(self, prefix, out)
        return out

    @staticmethod
    def _collect_state(unit: "FullyShardedDataParallel", prefix: str, out: Dict[str, torch.Tensor]) -> None:
        for k, v in unit._named_full_tensors(prefix):
            out[k] = v

        def visit(m: nn.Module, pfx: str) -> None:
            for bname, b in m._buffers.items():
                if b is not None and bname not in m._non_persistent_buffers_set:
                    out[pfx + bname] = b.detach().clone()
            for cname, c in m.named_children():
                if isinstance(c, FullyShardedDataParallel):
                    FullyShardedDataParallel._collect_state(c, pfx + cname + ".", out)
                else:
                    visit(c, pfx + cname + ".")

        visit(unit.module, prefix)

    def load_state_dict(self, state_dict: Dict[str, torch.Tensor], strict: bool = True):  # type: ignore[override]
        """Load a full (un-sharded) state dict; each rank keeps its slice."""
        missing: List[str] = []
        self._load_state(self, "", state_dict, missing)
        if strict and missing:
            raise KeyError(f"missing keys in state_dict: {missing[:8]}{'...' if len(missing) > 8 else ''}")
        return torch.nn.modules.module._IncompatibleKeys(missing, [])

    @staticmethod
    @torch.no_grad()
    def _load_state(unit: "FullyShardedDataParallel", prefix: str, sd: Dict[str, torch.Tensor], missing: List[str]) -> None:
        if unit.flat_param is not None:
            flat = torch.zeros(unit._padded, dtype=torch.float32, device=unit.flat_param.device)
            off = 0
            for (m, name, shape, n, fq) in unit._infos:
                key = prefix + fq
                if key in sd:
                    flat[off: off + n].copy_(sd[key].reshape(-1))
                else:
                    missing.append(key)
                off += n
            unit.flat_param.copy_(flat[unit.rank * unit._shard_numel: (unit.rank + 1) * unit._shard_numel])
            unit._full_valid = False

        def visit(m: nn.Module, pfx: str) -> None:
            for bname, b in m._buffers.items():
                if b is not None and bname not in m._non_persistent_buffers_set:
                    if pfx + bname in sd:
                        b.copy_(sd[pfx + bname])
                    else:
                        missing.append(pfx + bname)
            for cname, c in m.named_children():
                if isinstance(c, FullyShardedDataParallel):
                    FullyShardedDataParallel._load_state(c, pfx + cname + ".", sd, missing)
                else:
                    visit(c, pfx + cname + ".")

        visit(unit.module, prefix)

    def local_state_dict(self) -> Dict[str, torch.Tensor]:
        """Sharded state: one ``flat_param`` per unit, no communication."""
        out = {}
        for name, u in self.named_modules():
            if isinstance(u, FullyShardedDataParallel) and u.flat_param is not None:
                out[(name + "." if name else "") + "flat_param"] = u.flat_param.detach().clone()
        return out

    def load_local_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        with torch.no_grad():
            for name, u in self.named_modules():
                if isinstance(u, FullyShardedDataParallel) and u.flat_param is not None:
                    u.flat_param.copy_(sd[(name + "." if name else "") + "flat_param"])
                    u._full_valid = False

    def param_layout(self) -> List[Tuple[str, torch.Size, int, int]]:
        """(name, shape, offset, numel) of the original parameters inside the flat tensor."""
        out, off = [], 0
        for (_m, _n, shape, n, fq) in self._infos:
            out.append((fq, shape, off, n))
            off += n
        return out

    def gather_full_optim_state_dict(self, optim: torch.optim.Optimizer) -> Optional[Dict[str, Any]]:
        """Consolidate a sharded optimizer's state into per-original-parameter,
        un-flattened tensors on every rank (keys are fully-qualified names)."""
        from .optim_state import gather_full_optim_state
        return gather_full_optim_state(self, optim)

    def __getattr__(self, name: str) -> Any:
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self.module, name)


def default_auto_wrap_policy(module: nn.Module, unwrapped_params: int, min_num_params: int = int(1e6)) -> bool:
    return unwrapped_params >= min_num_params


def auto_wrap(module: nn.Module, policy: Optional[Callable[[nn.Module, int], bool]] = None,
              wrap_classes: Optional[Tuple[type, ...]] = None, **fsdp_kwargs: Any) -> nn.Module:
    """Wrap children bottom-up: a child becomes its own unit when it is an instance of
    ``wrap_classes`` or when ``policy(child, numel)`` says so.  The caller wraps the
    root."""
    policy = policy or default_auto_wrap_policy
    for name, child in list(module.named_children()):
        if isinstance(child, FullyShardedDataParallel):
            continue
        auto_wrap(child, policy, wrap_classes, **fsdp_kwargs)
        n = sum(p.numel() for p in FullyShardedDataParallel._own_params(child))
        hit = isinstance(child, wrap_classes) if wrap_classes else policy(child, n)
        if hit and n > 0:
            setattr(module, name, FullyShardedDataParallel(child, **fsdp_kwargs))
    return module

"""Gradient sharding (ZeRO-2): each gradient is reduced onto the one rank that owns
the parameter's optimizer state instead of being all-reduced everywhere.

Parity target: ``ShardedDDP(model, optimizer)`` (Fairscale-DDP.py:89) and the Stoke
flag ``fairscale_sddp=True`` (Stoke-DDP.py:250).  Constructor keywords follow
``FairscaleSDDPConfig``.

Mechanics
  * start-up: parameters and buffers are replicated from rank 0;
  * backward: a post-accumulate hook per parameter scales the gradient by
    1/world and issues ``reduce(dst=owner)``; parameters smaller than the bucket
    threshold share one flat bucket per destination rank, which is reduced when its
    last member fires.  Non-owners drop the gradient afterwards;
  * an end-of-backward callback waits on all outstanding handles;
  * ``no_sync()`` suppresses communication for gradient accumulation;
  * forward re-broadcasts buffers from rank 0 (``broadcast_buffers``).

On CUDA the facade attaches ``parallel.symm.ZeroPeerEngine``: the same hooks then
launch one peer-memory ``reduce_to_owner`` kernel per gradient bucket on a comm stream
while backward is still running, and ``step()`` finishes with the sharded AdamW + publish
kernel.  This module is the portable definition and the NCCL / gloo baseline.
"""
from __future__ import annotations

import contextlib
from typing import Any, Callable, Dict, Generator, List, Optional, Tuple, Union

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.autograd import Variable

from . import comm
from .oss import OSS

__all__ = ["ShardedDataParallel"]


class _Bucket:
    """Flat staging buffer for small gradients bound for one destination rank."""

    def __init__(self, size: int, dtype: torch.dtype, device: torch.device, dst: int):
        self.buffer = torch.zeros(size, dtype=dtype, device=device)
        self.dst = dst
        self.params: List[torch.Tensor] = []
        self.offsets: List[int] = []
        self.fill = 0
        self.pending = 0
        self.sent = False
        self.fired: set = set()      # ids of the members whose hook ran in this backward pass

    def can_add(self, p: torch.Tensor) -> bool:
        return self.fill + p.numel() <= self.buffer.numel()

    def add(self, p: torch.Tensor) -> None:
        self.params.append(p)
        self.offsets.append(self.fill)
        self.fill += p.numel()

    def reset(self) -> None:
        self.pending = len(self.params)
        self.sent = False
        self.fired.clear()


class ShardedDataParallel(nn.Module):
    def __init__(
        self,
        module: nn.Module,
        sharded_optimizer: Union[OSS, List[OSS]],
        process_group: Optional[Any] = None,
        broadcast_buffers: bool = True,
        sync_models_at_startup: bool = True,
        reduce_buffer_size: int = 2 ** 23,
        auto_refresh_trainable: bool = True,
        reduce_fp16: bool = False,
        warn_on_trainable_params_changed: bool = True,
    ):
        super().__init__()
        self.module = module
        self.sharded_optimizers: List[OSS] = (
            [sharded_optimizer] if not isinstance(sharded_optimizer, (list, tuple)) else list(sharded_optimizer)
        )
        for o in self.sharded_optimizers:
            if not isinstance(o, OSS):
                raise TypeError("ShardedDataParallel needs OSS-wrapped optimizer(s): gradient sharding "
                                "follows the optimizer's ownership map")
        self.process_group = process_group if process_group is not None else (
            dist.group.WORLD if comm.is_initialized() else None)
        self.world_size = comm.get_world_size(self.process_group)
        self.rank = comm.get_rank(self.process_group)
        self.broadcast_buffers = broadcast_buffers
        self.auto_refresh_trainable = auto_refresh_trainable
        self.reduce_fp16 = reduce_fp16
        self.reduce_buffer_size = reduce_buffer_size if self.world_size > 1 else 0
        self.should_accumulate_grads = False
        self.accumulate_grads_flipped = False
        self._fused_engine = None  # set by attach_fused_engine

        self._all_params = [p for p in module.parameters()]
        self._param_owner: Dict[torch.Tensor, Tuple[OSS, int]] = {}
        for o in self.sharded_optimizers:
            for p, r in o.param_to_rank.items():
                self._param_owner[p] = (o, r)
        self._trainable: List[torch.Tensor] = []
        self._trainable_mask: List[bool] = []
        self._hook_handles: List[Any] = []
        self._work: List[Tuple[Any, Optional[Callable[[], None]]]] = []
        self._buckets: Dict[Tuple[torch.device, torch.dtype, int], _Bucket] = {}
        self._param_bucket: Dict[torch.Tensor, Tuple[_Bucket, int]] = {}
        self._callback_queued = False

        self.refresh_trainable()
        if sync_models_at_startup:
            self.sync_params_and_buffers()

    # ------------------------------------------------------------------ set-up
    def _trainable_snapshot(self) -> List[bool]:
        return [p.requires_grad for p in self._all_params]

    def refresh_trainable(self) -> None:
        """Rebuild hooks and buckets after ``requires_grad`` changed."""
        self._trainable = [p for p in self._all_params if p.requires_grad]
        self._trainable_mask = self._trainable_snapshot()
        for p in self._trainable:
            if p not in self._param_owner:
                raise ValueError("a trainable parameter is not managed by any of the sharded optimizers")
        for o in self.sharded_optimizers:
            o.refresh_trainable()
        self._setup_buckets()
        self._setup_hooks()

    def _setup_buckets(self) -> None:
        self._buckets.clear()
        self._param_bucket.clear()
        if self.reduce_buffer_size <= 0 or self.world_size == 1:
            return
        for p in self._trainable:
            _, dst = self._param_owner[p]
            dtype = torch.float16 if (self.reduce_fp16 and p.dtype == torch.float32) else p.dtype
            key = (p.device, dtype, dst)
            b = self._buckets.get(key)
            if b is None:
                b = self._buckets[key] = _Bucket(self.reduce_buffer_size, dtype, p.device, dst)
            if b.can_add(p):
                self._param_bucket[p] = (b, b.fill)
                b.add(p)
        # trim, then arm
        for b in self._buckets.values():
            b.buffer = b.buffer[: max(b.fill, 1)].clone()
            b.reset()

    def _setup_hooks(self) -> None:
        for h in self._hook_handles:
            h.remove()
        self._hook_handles = []
        for p in self._trainable:
            self._hook_handles.append(p.register_post_accumulate_grad_hook(self._make_hook(p)))

    @torch.no_grad()
    def sync_params_and_buffers(self) -> None:
        """Replicate rank 0's parameters and buffers (K11, cold path)."""
        tensors = [p.data for p in self.module.parameters()] + [b.data for b in self.module.buffers()]
        comm.broadcast_coalesced(tensors, src=0, group=self.process_group)

    @torch.no_grad()
    def sync_buffers(self, blocking: bool = True) -> None:
        bufs = [b.data for b in self.module.buffers()]
        comm.broadcast_coalesced(bufs, src=0, group=self.process_group)

    def attach_fused_engine(self, engine: Any) -> None:
        self._fused_engine = engine

    # ------------------------------------------------------------------ forward
    def forward(self, *inputs: Any, **kwargs: Any) -> Any:
        if self.auto_refresh_trainable and self._trainable_snapshot() != self._trainable_mask:
            self.refresh_trainable()
        if self.broadcast_buffers and self.world_size > 1 and self.training and any(True for _ in self.module.buffers()):
            peer_sync = getattr(self._fused_engine, "sync_buffers", None)
            if peer_sync is not None:
                peer_sync()                      # one peer-memory kernel, no NCCL on the hot path
            else:
                self.sync_buffers()
        elif not self.training:
            self._flush_lazy_buffers()           # first eval forward after training: the statistics are read now
        if self._fused_engine is not None:
            self._fused_engine.pre_forward()
        for b in self._buckets.values():
            b.reset()
        self._callback_queued = False
        return self.module(*inputs, **kwargs)

    def to(self, *args: Any, **kwargs: Any) -> "ShardedDataParallel":  # type: ignore[override]
        self.module.to(*args, **kwargs)
        self.refresh_trainable()
        return self

    def zero_grad(self, set_to_none: bool = True) -> None:
        if self._fused_engine is not None:
            self._fused_engine.zero_grad()
            return
        for p in self._all_params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.detach_()
                    p.grad.zero_()

    @contextlib.contextmanager
    def no_sync(self) -> Generator:
        """Accumulate gradients locally; the next backward outside this context
        reduces the accumulated sum (Stoke micro-steps, Stoke-DDP.py:251)."""
        old = self.should_accumulate_grads
        self.should_accumulate_grads = True
        try:
            yield
        finally:
            self.should_accumulate_grads = old
            self.accumulate_grads_flipped = True

    # ------------------------------------------------------------------ backward
    def _queue_callback(self) -> None:
        if not self._callback_queued:
            self._callback_queued = True
            Variable._execution_engine.queue_callback(self._finalize_backward)

    def _make_hook(self, p: torch.Tensor) -> Callable[[torch.Tensor], None]:
        opt, dst = self._param_owner[p]
        gdst = comm.get_global_rank(self.process_group, dst)
        inv_world = 1.0 / self.world_size

        def hook(param: torch.Tensor) -> None:
            if self.should_accumulate_grads or self.world_size == 1:
                return
            if self._fused_engine is not None:
                # peer-memory path: the engine launches the bucket's reduce-to-owner on its comm stream as soon as
                # the bucket is complete; the end-of-backward callback joins the streams
                if self._fused_engine.grad_ready(param):
                    self._queue_callback()
                return
            grad = param.grad
            if grad is None:
                return
            self._queue_callback()
            entry = self._param_bucket.get(param)
            if entry is not None:
                bucket, off = entry
                view = bucket.buffer[off: off + param.numel()]
                # scale in the gradient's own precision first: a 16-bit bucket must not see the unscaled value
                view.copy_(grad.reshape(-1) * inv_world)
                bucket.fired.add(id(param))
                bucket.pending -= 1
                if bucket.pending == 0:
                    self._send_bucket(bucket)
                return
            # large parameter: reduce in place
            grad.mul_(inv_world)
            send = grad
            if self.reduce_fp16 and grad.dtype == torch.float32:
                send = grad.to(torch.float16)
            h = dist.reduce(send, dst=gdst, group=self.process_group, async_op=True)

            def done(param=param, send=send, dst=dst) -> None:
                if self.rank == dst:
                    if send is not param.grad:
                        param.grad.copy_(send)
                else:
                    param.grad = None

            self._work.append((h, done))

        return hook

    def _send_bucket(self, bucket: _Bucket) -> None:
        bucket.sent = True
        h = dist.reduce(bucket.buffer, dst=comm.get_global_rank(self.process_group, bucket.dst),
                        group=self.process_group, async_op=True)

        def done(bucket=bucket) -> None:
            if self.rank == bucket.dst:
                for p, off in zip(bucket.params, bucket.offsets):
                    if p.grad is not None:
                        p.grad.copy_(bucket.buffer[off: off + p.numel()].view_as(p))
            else:
                for p in bucket.params:
                    p.grad = None

        self._work.append((h, done))

    def _finalize_backward(self) -> None:
        """End-of-backward: flush partially filled buckets (unused parameters count
        as zero gradients), wait for every reduce, then drop non-owned gradients."""
        self._callback_queued = False
        if self._fused_engine is not None and not self.should_accumulate_grads:
            self._fused_engine.finish_backward()
        if not self.should_accumulate_grads and self.world_size > 1 and self._fused_engine is None:
            for b in self._buckets.values():
                if not b.sent and b.pending != len(b.params):
                    # some members never produced a gradient this pass: send zeros for them
                    for p, off in zip(b.params, b.offsets):
                        if id(p) not in b.fired:
                            b.buffer[off: off + p.numel()].zero_()
                    self._send_bucket(b)
        for h, done in self._work:
            h.wait()
            if done is not None:
                done()
        self._work.clear()
        self.accumulate_grads_flipped = False

    # convenience -------------------------------------------------------------
    def _flush_lazy_buffers(self) -> None:
        sync = getattr(self._fused_engine, "sync_buffers", None)
        if sync is not None and getattr(sync, "dirty", False) and self.world_size > 1:
            sync.flush()

    def state_dict(self, *args: Any, **kwargs: Any):  # type: ignore[override]
        self._flush_lazy_buffers()       # running statistics whose per-forward broadcast was elided (collective call)
        return self.module.state_dict(*args, **kwargs)

    def load_state_dict(self, *args: Any, **kwargs: Any):  # type: ignore[override]
        return self.module.load_state_dict(*args, **kwargs)

    def __getattr__(self, name: str) -> Any:
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self.module, name)

"""The single training object: owns model, optimizer, loss, scaler, gradient
accumulation, clipping, device placement, printing and checkpoint IO.

Parity target: ``Stoke(model=..., verbose=True, optimizer=StokeOptimizer(...), loss=...,
batch_size_per_device=..., gpu=True, fp16=None, distributed='ddp', fairscale_oss=True,
fairscale_sddp=True, grad_accum_steps=2, configs=[amp, ddp, oss],
grad_clip=ClipGradNormConfig(...))`` (Stoke-DDP.py:240-254) and every member the
script touches afterwards: ``.model`` ``.loss`` ``.backward`` ``.step`` (:73-82),
``.print_on_devices`` ``.print_ema_loss`` (:67,76,130), ``.model_access`` (:68,104),
``.detach_and_sync_loss`` (:86), ``.rank`` ``.world_size`` (:274-275), ``.DataLoader``
(:286-298), ``.optimizer`` (:300-301), ``.save`` (:142-145).

Build order: validate flags → process group → device placement →
SyncBN conversion → optimizer (OSS-wrapped when sharded) → engine wrap (DDP /
ShardedDDP / FSDP) → scaler.

B200-first choices
  * loss bookkeeping (last, accumulated, EMA) stays in device scalars; nothing on the
    training path calls ``.item()`` except the explicit print / sync methods;
  * on CUDA the optimizer class is swapped for the fused sm_100a implementation with
    identical state_dict layout, the scaler hands its inf flag to the update kernel as
    a device-side predicate, and the engines are attached to the peer-memory
    collectives when a symmetric-memory arena can be created.
"""
from __future__ import annotations

import contextlib
import os
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

import torch
import torch.distributed as dist
import torch.nn as nn

from . import io as _io
from .configs import (AMPConfig, B200Config, ClipGradConfig, ClipGradNormConfig, DDPConfig, DistributedOptions,
                      FairscaleFSDPConfig, FairscaleOSSConfig, FairscaleSDDPConfig, FP16Options, StokeOptimizer,
                      configs_to_dict)
from .data import StokeDataLoader
from .parallel import comm
from .parallel.ddp import DistributedDataParallel
from .parallel.fsdp import FullyShardedDataParallel, auto_wrap
from .parallel.grad_scaler import GradScaler, ShardedGradScaler
from .parallel.oss import OSS
from .parallel.sharded_ddp import ShardedDataParallel
from .parallel.sync_bn import convert_sync_batchnorm
from .status import RunStatus, resolve_status

__all__ = ["Stoke"]

LossT = Union[torch.Tensor, List[torch.Tensor], Tuple[torch.Tensor, ...]]


def _map_loss(fn: Callable[[torch.Tensor], Any], loss: LossT) -> Any:
    if isinstance(loss, (list, tuple)):
        return type(loss)(fn(l) for l in loss)
    return fn(loss)


class Stoke:
    def __init__(
        self,
        model: nn.Module,
        optimizer: StokeOptimizer,
        loss: Union[Callable, List[Callable], Tuple[Callable, ...]],
        batch_size_per_device: int,
        grad_accum_steps: Optional[int] = 1,
        grad_clip: Optional[Union[ClipGradConfig, ClipGradNormConfig]] = None,
        gpu: bool = False,
        fp16: Optional[Union[str, FP16Options]] = None,
        distributed: Optional[Union[str, DistributedOptions]] = None,
        fairscale_oss: bool = False,
        fairscale_sddp: bool = False,
        fairscale_fsdp: bool = False,
        configs: Optional[List[Any]] = None,
        info_rank: Optional[Union[int, List[int]]] = 0,
        verbose: bool = True,
        ema_weight: float = 0.1,
    ):
        if gpu and not torch.cuda.is_available() and os.environ.get("B200DP_CPU_FALLBACK", "0") not in ("", "0"):
            # scripts written for a GPU box (Stoke-DDP.py:246 hard-codes gpu=True) can be exercised on a CPU-only
            # machine: everything runs through the portable engines over gloo
            print("Stoke -- gpu=True but no CUDA device is visible; B200DP_CPU_FALLBACK is set: running on the CPU", flush=True)
            gpu = False
        self._verbose = verbose
        self._info_rank = info_rank
        self._ema_weight = float(ema_weight)
        self._configs = configs_to_dict(configs)
        self._status: RunStatus = resolve_status(batch_size_per_device, grad_accum_steps, gpu, fp16, distributed,
                                                 fairscale_oss, fairscale_sddp, fairscale_fsdp)
        if grad_clip is not None and not isinstance(grad_clip, (ClipGradConfig, ClipGradNormConfig)):
            raise TypeError("grad_clip must be a ClipGradConfig or ClipGradNormConfig")
        self._grad_clip = grad_clip
        self._status.grad_clip = None if grad_clip is None else {"type": type(grad_clip).__name__, **vars(grad_clip)}
        self._b200: B200Config = self._configs.get(B200Config, B200Config())
        self._amp_cfg: AMPConfig = self._configs.get(AMPConfig, AMPConfig())
        self._ddp_cfg: DDPConfig = self._configs.get(DDPConfig, DDPConfig())
        self._oss_cfg: FairscaleOSSConfig = self._configs.get(FairscaleOSSConfig, FairscaleOSSConfig())
        self._sddp_cfg: FairscaleSDDPConfig = self._configs.get(FairscaleSDDPConfig, FairscaleSDDPConfig())
        self._fsdp_cfg: FairscaleFSDPConfig = self._configs.get(FairscaleFSDPConfig, FairscaleFSDPConfig())

        # counters (checkpointed)
        self._backward_steps = 0
        self._optimizer_steps = 0
        # host-side hang guard (utils.watchdog): beats once per train_step / optimizer step
        self._watchdog = None
        if self._b200.step_watchdog_s > 0:
            from .utils.watchdog import StepWatchdog
            self._watchdog = StepWatchdog(self._b200.step_watchdog_s, abort=self._b200.step_watchdog_abort)
        self._grad_accum_counter = 0
        # loss tracking, device scalars
        self._last_step_loss: Any = None
        self._agg_loss: Any = None
        self._ema_loss: Any = None
        self._loss_sync_calls = 0
        self._skipped_optimizer_steps = 0
        self._agg_fresh = True
        self._graphs: Dict[Any, Dict[str, Any]] = {}

        # 1. process group + device
        self._device = self._setup_device_and_group(gpu)
        self._status.world_size = self.world_size
        # 2. model placement
        self._loss = loss
        self._raw_model = model
        if gpu:
            model = model.to(self._device)
            if self._b200.channels_last and any(p.dim() == 4 for p in model.parameters()):
                model = model.to(memory_format=torch.channels_last)
        if self._status.is_distributed and self._ddp_cfg.convert_to_sync_batch_norm and self.world_size > 1:
            model = convert_sync_batchnorm(model)
        if gpu and self._b200.fused_kernels:
            # conv -> BatchNorm pairs: the convolution's epilogue emits the batch statistics
            from .ops.layers import pair_conv_bn
            pair_conv_bn(model)
        self._unwrapped = model
        # 3/4. optimizer + engine
        self._optimizer_spec = optimizer
        self._model, self._optimizer = self._build_engine(model, optimizer)
        # 16-bit weight shadows: the fused update kernel also writes the compute-dtype copy of every matrix-shaped
        # parameter, so convolutions / linears stop casting their fp32 weights each forward
        if self._fused_ok() and self._status.fp16 is not None and not self._status.fully_sharded:
            inner = self._optimizer.optim if isinstance(self._optimizer, OSS) else self._optimizer
            if getattr(inner, "writes_shadows", False) and getattr(self._optimizer, "_flat_engine", None) is None:
                inner.lp_dtype = torch.bfloat16 if self._status.is_bf16 else torch.float16
                inner.make_shadows()
        # 5. scaler
        self._scaler = self._build_scaler()
        if self._verbose:
            self.print(self._describe())

    # ------------------------------------------------------------------ construction
    def _setup_device_and_group(self, gpu: bool) -> torch.device:
        local_rank = self._ddp_cfg.local_rank
        if local_rank is None:
            env = os.environ.get("LOCAL_RANK")
            local_rank = int(env) if env not in (None, "None", "") else 0
        device = torch.device("cpu")
        if gpu:
            if not torch.cuda.is_available():
                raise RuntimeError("gpu=True but no CUDA device is visible")
            device = torch.device("cuda", int(local_rank) % torch.cuda.device_count())
            torch.cuda.set_device(device)
        if self._status.is_distributed:
            backend = self._ddp_cfg.backend
            if backend == "nccl" and not gpu:
                backend = "gloo"
            comm.init_process_group(backend=backend, init_method=self._ddp_cfg.init_method,
                                    timeout_s=max(self._b200.watchdog_timeout_s * 10, 600.0),
                                    device_id=device if gpu else None)
        return device

    def _fused_ok(self) -> bool:
        if not (self._status.gpu and self._b200.fused_kernels):
            return False
        from .ops import library_available, require_library
        if self._b200.strict:
            require_library()
        return library_available()

    def _optimizer_class(self, cls: type) -> type:
        """Swap stock optimizers for the fused sm_100a ones on CUDA (same hyper-parameters,
        same state_dict layout)."""
        if not self._fused_ok():
            return cls
        from .ops import optim as fused_optim
        return fused_optim.fused_equivalent(cls)

    def _build_engine(self, model: nn.Module, spec: StokeOptimizer) -> Tuple[nn.Module, torch.optim.Optimizer]:
        st = self._status
        opt_cls = self._optimizer_class(spec["optimizer"])
        opt_kwargs = dict(spec.get("optimizer_kwargs", {}))
        if st.fully_sharded:
            c = self._fsdp_cfg
            compute_dtype = c.compute_dtype
            if compute_dtype is None and st.fp16 is not None:
                compute_dtype = torch.bfloat16 if st.is_bf16 else torch.float16
            kw = dict(reshard_after_forward=c.reshard_after_forward, mixed_precision=c.mixed_precision or st.fp16 is not None,
                      fp32_reduce_scatter=c.fp32_reduce_scatter, flatten_parameters=c.flatten_parameters,
                      compute_dtype=compute_dtype, buffer_dtype=c.buffer_dtype, bucket_cap_mb=c.bucket_cap_mb,
                      force_input_to_fp32=c.force_input_to_fp32, verbose=c.verbose,
                      gradient_predivide_factor=c.gradient_predivide_factor,
                      gradient_postdivide_factor=c.gradient_postdivide_factor)
            wrap_cls = getattr(model, "fsdp_wrap_classes", None)
            model = auto_wrap(model, policy=lambda m, n: n >= c.min_num_params, wrap_classes=wrap_cls, **kw)
            engine = FullyShardedDataParallel(model, **kw)
            optim = opt_cls(engine.parameters(), **opt_kwargs)
            self._attach_peer(engine, optim)
            return engine, optim
        if st.oss:
            optim = OSS(model.parameters(), optim=opt_cls, broadcast_fp16=self._oss_cfg.broadcast_fp16, **opt_kwargs)
            if st.sharded:
                c = self._sddp_cfg
                engine: nn.Module = ShardedDataParallel(
                    model, optim, broadcast_buffers=c.broadcast_buffers,
                    sync_models_at_startup=c.sync_models_at_startup, reduce_buffer_size=c.reduce_buffer_size,
                    auto_refresh_trainable=c.auto_refresh_trainable, reduce_fp16=c.reduce_fp16)
            else:
                engine = self._wrap_ddp(model)
            self._attach_peer(engine, optim)
            return engine, optim
        if st.is_distributed:
            engine = self._wrap_ddp(model)
            optim = opt_cls(model.parameters(), **opt_kwargs)
            self._attach_peer(engine, optim)
            return engine, optim
        if self._b200.cuda_graph and st.gpu:
            # graph capture needs gradient addresses that never change: alias them to flat buckets
            engine = DistributedDataParallel(model, sync_models_at_startup=False)
            return engine, opt_cls(model.parameters(), **opt_kwargs)
        return model, opt_cls(model.parameters(), **opt_kwargs)

    def _wrap_ddp(self, model: nn.Module) -> DistributedDataParallel:
        c = self._ddp_cfg
        return DistributedDataParallel(model, broadcast_buffers=c.broadcast_buffers, bucket_cap_mb=c.bucket_cap_mb,
                                       find_unused_parameters=c.find_unused_parameters)

    def _attach_peer(self, engine: nn.Module, optim: torch.optim.Optimizer) -> None:
        """Attach the peer-memory fused collectives when possible (CUDA, >1 rank, kernel
        library built, symmetric memory available).  Otherwise the engines keep their
        ``torch.distributed`` collectives — the measured baseline."""
        self._peer = None
        if not (self._fused_ok() and self._b200.peer_collectives and self.world_size > 1):
            return
        try:
            from .parallel import peer
            self._peer = peer.attach(engine, optim, self._status, scaler_cfg=self._amp_cfg, ddp_cfg=self._ddp_cfg,
                                     oss_cfg=self._oss_cfg, timeout_s=self._b200.watchdog_timeout_s)
        except Exception as e:  # pragma: no cover - depends on the box
            if self._b200.strict:
                raise RuntimeError(f"peer-memory collectives could not be attached (B200Config.strict): {e}") from e
            if self._verbose:
                self.print(f"[b200] peer-memory collectives unavailable ({type(e).__name__}: {e}); using NCCL")
            self._peer = None

    def _build_scaler(self) -> Optional[GradScaler]:
        if not self._status.is_amp:
            return None
        c = self._amp_cfg
        kw = dict(init_scale=c.init_scale, growth_factor=c.growth_factor, backoff_factor=c.backoff_factor,
                  growth_interval=c.growth_interval, enabled=c.enabled)
        if self._status.is_fairscale and self.world_size > 1:
            return ShardedGradScaler(**kw)
        return GradScaler(**kw)

    def _describe(self) -> str:
        s = self._status
        eng = type(self._model).__name__ if self._model is not self._unwrapped else "none (single process)"
        return (f"Stoke -- device={self._device} world_size={self.world_size} engine={eng} "
                f"optimizer={type(self._optimizer).__name__}"
                f"{'(' + type(self._optimizer.optim).__name__ + ')' if isinstance(self._optimizer, OSS) else ''} "
                f"fp16={s.fp16} grad_accum={s.grad_accum_steps} batch/device={s.batch_size_per_device} "
                f"effective_batch={s.effective_batch_size} clip={s.grad_clip} "
                f"fused_kernels={self._fused_ok()} peer_collectives={getattr(self, '_peer', None) is not None}")

    # ------------------------------------------------------------------ forward / loss
    def _autocast(self):
        if self._status.fp16 is None or not self._status.gpu:
            return contextlib.nullcontext()
        dtype = torch.bfloat16 if self._status.is_bf16 else torch.float16
        return torch.autocast(device_type="cuda", dtype=dtype)

    def model(self, *args: Any, **kwargs: Any) -> Any:
        """Forward through the wrapped engine under the precision context (Stoke-DDP.py:73)."""
        with self._autocast():
            return self._model(*args, **kwargs)

    def loss(self, *args: Any, **kwargs: Any) -> LossT:
        """Loss under the precision context (Stoke-DDP.py:74,118).

        Training mode: records the step loss, accumulates it, updates the EMA and
        returns ``loss / grad_accum_steps`` (ready for ``backward``).  Eval mode: returns
        the raw loss and touches no bookkeeping (the validation loop relies on that).
        """
        with self._autocast():
            if isinstance(self._loss, (list, tuple)):
                loss = type(self._loss)(fn(*args, **kwargs) for fn in self._loss)
            else:
                loss = self._loss(*args, **kwargs)
        if not self._unwrapped.training:
            return loss
        w = self._ema_weight
        if isinstance(loss, torch.Tensor):
            # persistent device scalars updated in place: no per-step allocations, no host reads, and
            # a CUDA-graph replay keeps feeding the same buffers
            det = loss.detach().float()
            if self._last_step_loss is None or not isinstance(self._last_step_loss, torch.Tensor):
                self._last_step_loss, self._agg_loss, self._ema_loss = det.clone(), det.clone(), det.clone()
                self._agg_fresh = False
            else:
                self._last_step_loss.copy_(det)
                if self._agg_loss is None:
                    self._agg_loss = det.clone()
                elif self._agg_fresh:
                    self._agg_loss.copy_(det)
                else:
                    self._agg_loss.add_(det)
                self._agg_fresh = False
                if self._ema_loss is None:
                    self._ema_loss = det.clone()
                else:
                    self._ema_loss.mul_(1.0 - w).add_(det, alpha=w)
        else:
            det = _map_loss(lambda l: l.detach().float(), loss)
            self._last_step_loss = det
            self._agg_loss = det if (self._agg_loss is None or self._agg_fresh) else self._zip(lambda a, b: a + b, self._agg_loss, det)
            self._agg_fresh = False
            self._ema_loss = det if self._ema_loss is None else self._zip(lambda e, l: e * (1.0 - w) + l * w,
                                                                           self._ema_loss, det)
        if self.grad_accum > 1:
            return _map_loss(lambda l: l / self.grad_accum, loss)
        return loss

    @staticmethod
    def _zip(fn: Callable, a: Any, b: Any) -> Any:
        if isinstance(a, (list, tuple)):
            return type(a)(fn(x, y) for x, y in zip(a, b))
        return fn(a, b)

    # ------------------------------------------------------------------ backward / step
    def _at_boundary(self) -> bool:
        return self._grad_accum_counter == self.grad_accum

    def backward(self, loss: LossT) -> None:
        """Backward with accumulation-aware communication (Stoke-DDP.py:79): on
        non-boundary micro-steps the engine's ``no_sync()`` suppresses the collective."""
        self._backward_steps += 1
        self._grad_accum_counter += 1
        total = sum(loss) if isinstance(loss, (list, tuple)) else loss
        sync_now = self._at_boundary()
        if self._scaler is not None:
            total = self._scaler.scale(total)
            ddp_eng = self._peer.get("ddp") if getattr(self, "_peer", None) else None
            if ddp_eng is not None and sync_now:
                # the fused bucket all-reduce applies 1/scale and raises the inf flag itself (W1)
                if ddp_eng.inv_scale is None:
                    ddp_eng.inv_scale = torch.empty((), dtype=torch.float32, device=self._device)
                torch.reciprocal(self._scaler.scale_tensor, out=ddp_eng.inv_scale)
                ddp_eng.found_inf = self._scaler.found_inf_tensor
            zero_eng = getattr(self._optimizer, "_flat_engine", None) if isinstance(self._optimizer, OSS) else None
            if zero_eng is not None and sync_now:
                # the bucket reduce-to-owner kernels of this backward apply 1/scale themselves (W2)
                if zero_eng.inv_scale is None:
                    zero_eng.inv_scale = torch.empty((), dtype=torch.float32, device=self._device)
                torch.reciprocal(self._scaler.scale_tensor, out=zero_eng.inv_scale)
        ctx = contextlib.nullcontext()
        if not sync_now and self._ddp_cfg.no_sync and hasattr(self._model, "no_sync"):
            ctx = self._model.no_sync()
        with ctx:
            total.backward()

    def step(self) -> None:
        """Optimizer step at accumulation boundaries, a no-op otherwise (Stoke-DDP.py:82):
        (unscale) → clip → step (skipped on inf) → scaler update → zero grads."""
        if not self._at_boundary():
            return
        gc = self._grad_clip
        opt = self._optimizer
        norm_clip = isinstance(gc, ClipGradNormConfig) and float(gc.norm_type) == 2.0
        fused = hasattr(opt, "fused_unscale_clip") and not isinstance(gc, ClipGradConfig) and (gc is None or norm_clip)
        if fused and isinstance(opt, OSS) and not opt._fused_clip_usable():
            fused = False
        zero_eng = getattr(opt, "_flat_engine", None) if isinstance(opt, OSS) else None
        ddp_eng = self._peer.get("ddp") if getattr(self, "_peer", None) else None
        if zero_eng is not None and (gc is None or norm_clip):
            # fused ZeRO-2 step: reduce-to-owner (+norm, +inf) -> tiny all-reduce -> AdamW + publish
            if self._scaler is not None:
                self._scaler._lazy_init(self._device)
                inv = self._scaler.scale_tensor.reciprocal()
                zero_eng.clip_grad_norm(gc.max_norm if gc is not None else 3.0e38, 2.0, inv_scale=inv,
                                        found_inf=self._scaler.found_inf_tensor)
                self._scaler.unscale_(opt, already_unscaled=True, flag_is_global=True)
                self._scaler.step(opt)
                self._scaler.update()
            else:
                if gc is not None:
                    zero_eng.clip_grad_norm(gc.max_norm, 2.0)
                opt.step()
        elif self._scaler is not None:
            if ddp_eng is not None and ddp_eng.amp_fused:
                # gradients arrive from the fused all-reduce already unscaled, flag already raised
                self._scaler.unscale_(opt, already_unscaled=True, flag_is_global=True)
                if fused and gc is not None:
                    opt.fused_unscale_clip(None, gc.max_norm, self._scaler.found_inf_tensor)
                else:
                    self._clip()
            elif fused:
                # one read of the gradients (inf flag + norm); unscale and clip ride inside the update kernel
                self._scaler.unscale_and_clip_fused(opt, None if gc is None else gc.max_norm)
            else:
                if gc is not None:
                    self._scaler.unscale_(opt)
                self._clip()
            self._scaler.step(opt)
            self._scaler.update()
        else:
            self._clip()
            opt.step()
        self.zero_grads()
        self._grad_accum_counter = 0
        self._agg_fresh = True          # the next training loss restarts the accumulation window
        self._optimizer_steps += 1
        from .ops.fused import bump_param_epoch
        bump_param_epoch()              # value caches of parameters (zero-padded copies) are stale now
        if self._watchdog is not None:
            self._watchdog.beat()

    # ------------------------------------------------------------------ fused train step (+ CUDA graph)
    def train_step(self, inputs: Any, targets: Any) -> Any:
        """One micro-step as a single call: ``loss = self.loss(self.model(inputs), targets)``,
        ``self.backward(loss)``, ``self.step()``.  Returns the detached (undivided) loss scalar,
        a device tensor.

        With ``B200Config(cuda_graph=True)`` on a GPU the micro-step is captured into a CUDA graph
        after three eager warm-up calls and replayed afterwards: one launch instead of ~1000, so the
        step stops being bound by Python / launch overhead.  Everything the step reads on the host
        is baked in at capture time, except the learning rate (kept in a device scalar that is
        refreshed before every replay, so LR schedulers keep working) — so change other
        hyper-parameters, ``requires_grad`` flags or input shapes only via a new shape (new graph) or
        ``reset_graphs()``.  Inputs are copied into static buffers before each replay.
        """
        args = inputs if isinstance(inputs, (tuple, list)) else (inputs,)
        if self._watchdog is not None:
            self._watchdog.beat()
        if not (self._b200.cuda_graph and self._status.gpu and self._unwrapped.training):
            return self._eager_step(args, targets)
        return self._graphed_step(args, targets)

    def _eager_step(self, args: Any, targets: Any) -> Any:
        loss = self.loss(self.model(*args), targets)
        self.backward(loss)
        self.step()
        return self._last_step_loss

    def reset_graphs(self) -> None:
        self._graphs.clear()

    def _sync_lr(self) -> None:
        opt = self._optimizer.optim if isinstance(self._optimizer, OSS) else self._optimizer
        if hasattr(opt, "sync_lr"):
            if isinstance(self._optimizer, OSS):
                self._optimizer._sync_param_groups(self._optimizer.param_groups, opt.param_groups)
            opt.sync_lr()

    @property
    def native_status(self) -> Dict[str, Any]:
        """What is really serving this run: kernel library, peer-memory engines, CUDA graph."""
        from .ops import _lib
        peer = getattr(self, "_peer", None)
        return {"fused_kernels": self._fused_ok(), "kernel_library": _lib.lib_path() if self._fused_ok() else None,
                "peer_collectives": sorted(k for k in peer if k not in ("arena", "small")) if peer else [],
                "nvls_multicast": bool(peer and peer["arena"].mc_base),
                "cuda_graph": bool(self._b200.cuda_graph and any(e.get("graph") is not None for e in self._graphs.values()))}

    def _graphed_step(self, args: Any, targets: Any) -> Any:
        flat = [a for a in args if isinstance(a, torch.Tensor)] + ([targets] if isinstance(targets, torch.Tensor) else [])
        key = (self._grad_accum_counter, tuple((tuple(t.shape), t.dtype, t.stride()) for t in flat))
        ent = self._graphs.setdefault(key, {"calls": 0, "graph": None})
        if ent["graph"] is None:
            if ent["calls"] < 3:
                ent["calls"] += 1
                return self._eager_step(args, targets)
            # ---- capture
            ent["args"] = tuple(a.clone() if isinstance(a, torch.Tensor) else a for a in args)
            ent["targets"] = targets.clone() if isinstance(targets, torch.Tensor) else targets
            self._sync_lr()
            counters = (self._backward_steps, self._optimizer_steps, self._grad_accum_counter, self._agg_fresh)
            torch.cuda.synchronize(self._device)
            graph = torch.cuda.CUDAGraph()
            from .ops import _lib
            launches0 = _lib.LAUNCHES
            try:
                with torch.cuda.graph(graph):
                    self._eager_step(ent["args"], ent["targets"])
            except Exception as e:  # this configuration cannot be captured: stay eager from here on
                self._backward_steps, self._optimizer_steps, self._grad_accum_counter, self._agg_fresh = counters
                self._b200.cuda_graph = False
                self._graphs.clear()
                torch.cuda.synchronize(self._device)
                self.reset()
                if self._b200.strict:
                    raise
                self.print(f"[b200] CUDA-graph capture failed ({type(e).__name__}: {str(e)[:160]}); continuing eagerly")
                return self._eager_step(args, targets)
            ent["launches"] = _lib.LAUNCHES - launches0    # b200dp kernels inside one replay
            _lib.LAUNCHES = launches0
            ent["delta"] = (self._backward_steps - counters[0], self._optimizer_steps - counters[1],
                            self._grad_accum_counter, self._agg_fresh)
            ent["graph"] = graph
            # capture records, it does not execute: roll the host counters back and replay for real
            self._backward_steps, self._optimizer_steps, self._grad_accum_counter, self._agg_fresh = counters
        for dst, src in zip(ent["args"], args):
            if isinstance(dst, torch.Tensor):
                dst.copy_(src, non_blocking=True)
        if isinstance(ent["targets"], torch.Tensor):
            ent["targets"].copy_(targets, non_blocking=True)
        self._sync_lr()
        ent["graph"].replay()
        from .ops import _lib
        _lib.LAUNCHES += ent["launches"]
        d = ent["delta"]
        self._backward_steps += d[0]
        self._optimizer_steps += d[1]
        self._grad_accum_counter, self._agg_fresh = d[2], d[3]
        if d[1]:
            from .ops.fused import bump_param_epoch
            bump_param_epoch()          # the replay updated the parameters without running any Python
        return self._last_step_loss

    def _clip(self) -> None:
        """Clip dispatch."""
        gc = self._grad_clip
        if gc is None:
            return
        if isinstance(gc, ClipGradConfig):
            params = [p for g in self._optimizer.param_groups for p in g["params"] if p.grad is not None]
            torch.nn.utils.clip_grad_value_(params, gc.clip_value)
            return
        if isinstance(self._optimizer, OSS):
            self._optimizer.clip_grad_norm(gc.max_norm, gc.norm_type)
        elif isinstance(self._model, FullyShardedDataParallel):
            self._model.clip_grad_norm_(gc.max_norm, gc.norm_type)
        elif hasattr(self._optimizer, "clip_grad_norm"):
            self._optimizer.clip_grad_norm(gc.max_norm, gc.norm_type)
        else:
            params = [p for g in self._optimizer.param_groups for p in g["params"] if p.grad is not None]
            torch.nn.utils.clip_grad_norm_(params, gc.max_norm, gc.norm_type)

    def zero_grads(self) -> None:
        if isinstance(self._model, DistributedDataParallel):
            self._model.zero_grad()
        elif isinstance(self._model, ShardedDataParallel):
            self._model.zero_grad(set_to_none=not (self._b200.cuda_graph and self._status.gpu))
        else:
            # under CUDA graphs gradient tensors must keep their addresses: zero in place
            self._optimizer.zero_grad(set_to_none=not (self._b200.cuda_graph and self._status.gpu))

    def reset(self) -> None:
        """Drop gradients and accumulation state (e.g. after a manual interruption)."""
        self.zero_grads()
        self._grad_accum_counter = 0
        self._agg_fresh = True

    def reset_ema(self) -> None:
        self._ema_loss = None

    def reset_tracking(self) -> None:
        self._last_step_loss = self._agg_loss = self._ema_loss = None

    # ------------------------------------------------------------------ loss sync / printing
    def detach_and_sync_loss(self, loss: LossT, device: Optional[torch.device] = None) -> Any:
        """Mean of a loss across ranks as a Python float (Stoke-DDP.py:86).

        The reference pays one NCCL all-reduce and one blocking ``.item()`` per batch here.  On the GPU path the
        cross-rank sum is the one-shot peer-memory kernel (W11: no NCCL launch), the scalar is copied into pinned
        host memory asynchronously, and
          * ``B200Config(loss_sync_lag=0)`` (default) waits for that copy: the reference's exact semantics;
          * ``loss_sync_lag=1`` returns the value of the PREVIOUS call instead (first call: this one's), so the host
            never waits for the step it has just enqueued — the loss curve is the same, one call late;
          * ``loss_sync_every=k`` refreshes only every k-th call and returns the last value in between.
        """
        self._loss_sync_calls += 1
        k = max(1, self._b200.loss_sync_every)
        if k > 1 and (self._loss_sync_calls % k) != 0 and getattr(self, "_last_synced", None) is not None:
            return self._last_synced
        dist_on = self._status.is_distributed and self.world_size > 1

        def one_host(l: torch.Tensor) -> float:
            t = l.detach().float().clone()
            if dist_on:
                dist.all_reduce(t)
                t /= self.world_size
            return float(t.item())

        losses = list(loss) if isinstance(loss, (list, tuple)) else [loss]
        on_gpu = self._status.gpu and all(isinstance(l, torch.Tensor) and l.is_cuda for l in losses)
        if not on_gpu:
            self._last_synced = _map_loss(one_host, loss)
            return self._last_synced
        # ---- device path: pack, one small all-reduce, async drain
        n = len(losses)
        st = getattr(self, "_loss_sync_state", None)
        if st is None or st["dev"].numel() != n:
            st = self._loss_sync_state = {
                "dev": torch.zeros(n, dtype=torch.float32, device=self._device),
                "host": [torch.zeros(n, dtype=torch.float32).pin_memory() for _ in range(2)],
                "evt": [torch.cuda.Event(), torch.cuda.Event()], "slot": 0, "pending": None}
        buf = st["dev"]
        for i, l in enumerate(losses):
            buf[i].copy_(l.detach().float().reshape(()))
        if dist_on:
            from .parallel import peer
            if not peer.small_allreduce(buf, None):
                dist.all_reduce(buf)
            buf.div_(self.world_size)
        slot = st["slot"]
        st["host"][slot].copy_(buf, non_blocking=True)
        st["evt"][slot].record()
        lag = max(0, int(getattr(self._b200, "loss_sync_lag", 0)))
        read = slot if (lag == 0 or st["pending"] is None) else st["pending"]
        st["evt"][read].synchronize()
        vals = [float(v) for v in st["host"][read]]
        st["pending"], st["slot"] = slot, slot ^ 1
        self._last_synced = type(loss)(vals) if isinstance(loss, (list, tuple)) else vals[0]
        return self._last_synced

    def _is_info_rank(self) -> bool:
        ir = self._info_rank
        if ir is None:
            return True
        if isinstance(ir, (list, tuple)):
            return self.rank in ir
        return self.rank == ir

    def print(self, msg: Union[str, List[str]], single_line: bool = False) -> None:
        if self._is_info_rank():
            if isinstance(msg, (list, tuple)):
                msg = (" ".join(msg)) if single_line else "\n".join(msg)
            print(f"Stoke -- {msg}" if not str(msg).startswith("Stoke --") else msg, flush=True)

    def print_on_devices(self, msg: Union[str, List[str]], rank: Optional[Union[int, List[int]]] = 0) -> None:
        """Rank-filtered print (Stoke-DDP.py:67,130)."""
        ranks = None if rank is None else ([rank] if isinstance(rank, int) else list(rank))
        if ranks is None or self.rank in ranks:
            if isinstance(msg, (list, tuple)):
                msg = "\n".join(msg)
            print(f"Stoke -- {msg}", flush=True)

    def print_ema_loss(self, prepend_msg: str = "Current EMA Loss", single_line: bool = False) -> None:
        """Stoke-DDP.py:76.  Reads a device scalar (host sync) only on the info rank."""
        if self._ema_loss is None or not self._is_info_rank() or not self._verbose:
            return
        vals = _map_loss(lambda t: float(t.item()), self._ema_loss)
        self.print(f"{prepend_msg}: " + (", ".join(f"{v:.3f}" for v in vals) if isinstance(vals, (list, tuple)) else f"{vals:.3f}"))

    def print_mean_accumulated_synced_loss(self, prepend_msg: str = "Mean Accumulated & Synced Loss") -> None:
        if self._agg_loss is None or self._agg_fresh:
            return
        n = max(self._grad_accum_counter, 1)
        vals = self.detach_and_sync_loss(_map_loss(lambda t: t / n, self._agg_loss))
        self.print(f"{prepend_msg}: {vals}")

    def print_synced_loss(self, loss: LossT, prepend_msg: str = "Step Synced Loss", scale_factor: float = 1.0) -> None:
        vals = self.detach_and_sync_loss(_map_loss(lambda t: t * scale_factor, loss))
        self.print(f"{prepend_msg}: {vals}")

    def print_num_model_parameters(self, normalize: str = "M") -> None:
        div = {"": 1.0, "K": 1e3, "M": 1e6, "B": 1e9}.get(normalize, 1e6)
        self.print(f"Current Model Parameters: {self.num_model_parameters / div:.3f} {normalize}")

    def dump_model_parameter_info(self) -> None:
        self.print("Dumping all model parameter information to stdout....")
        for name, p in self._unwrapped.named_parameters():
            if p.requires_grad:
                self.print(f"Name: {name}, Shape: {tuple(p.shape)}, Parameters: {p.numel()}")

    def barrier(self) -> None:
        comm.barrier()

    # ------------------------------------------------------------------ data
    def DataLoader(self, dataset: Any, **kwargs: Any) -> StokeDataLoader:  # noqa: N802 (reference name)
        """Loader bound to this object's batch size and device (Stoke-DDP.py:286-298)."""
        if "batch_size" in kwargs:
            raise ValueError("batch_size is set by the Stoke object (batch_size_per_device); do not pass it")
        return StokeDataLoader(dataset, batch_size=self.batch_size, gpu=self._status.gpu, device=self._device,
                               distributed=self._status.is_distributed and self.world_size > 1,
                               channels_last=self._b200.channels_last and self._status.gpu, **kwargs)

    # ------------------------------------------------------------------ IO
    def save(self, path: str, name: str = "", extension: str = "pt", create_directory: bool = True,
             extras: Optional[dict] = None) -> Tuple[str, str]:
        """Checkpoint model / optimizer / scaler / counters (Stoke-DDP.py:142-145).
        Collective: every rank must call it."""
        return _io.save(self, path=path, name=name, extension=extension, create_directory=create_directory,
                        extras=extras)

    def load(self, path: str, tag: str, strict: bool = True) -> Optional[dict]:
        """Resume from a checkpoint written by :meth:`save`; returns ``extras``."""
        return _io.load(self, path=path, tag=tag, strict=strict)

    # ------------------------------------------------------------------ accessors
    @property
    def model_access(self) -> nn.Module:
        """The bare ``nn.Module`` (``.train()`` / ``.eval()``; Stoke-DDP.py:68,104)."""
        return self._unwrapped

    @property
    def wrapped_model(self) -> nn.Module:
        return self._model

    @property
    def loss_access(self) -> Any:
        return self._loss

    @property
    def optimizer(self) -> torch.optim.Optimizer:
        """Live optimizer; exposes ``param_groups`` for LR schedulers (Stoke-DDP.py:300-306)."""
        return self._optimizer

    @property
    def scaler(self) -> Optional[GradScaler]:
        return self._scaler

    @property
    def rank(self) -> int:
        return comm.get_rank() if self._status.is_distributed else 0

    @property
    def world_size(self) -> int:
        return comm.get_world_size() if self._status.is_distributed else 1

    @property
    def device(self) -> torch.device:
        return self._device

    @property
    def batch_size(self) -> int:
        return self._status.batch_size_per_device

    @property
    def effective_batch_size(self) -> int:
        return self._status.effective_batch_size

    @property
    def grad_accum(self) -> int:
        return self._status.grad_accum_steps

    @property
    def grad_clip(self) -> Optional[Union[ClipGradConfig, ClipGradNormConfig]]:
        return self._grad_clip

    @property
    def ema_loss(self) -> Any:
        return self._ema_loss

    @property
    def step_loss(self) -> Any:
        return self._last_step_loss

    @property
    def num_model_parameters(self) -> int:
        if isinstance(self._model, FullyShardedDataParallel):
            return sum(u._total for u in self._model._fsdp_units())
        return sum(p.numel() for p in self._unwrapped.parameters() if p.requires_grad)

    @property
    def status(self) -> Dict[str, Any]:
        return self._status.to_dict()

    @property
    def backward_steps(self) -> int:
        return self._backward_steps

    @property
    def optimizer_steps(self) -> int:
        return self._optimizer_steps

    @property
    def watchdog(self) -> Any:
        """The host-side ``StepWatchdog`` (``B200Config.step_watchdog_s > 0``) or None; ``pause()`` it around
        legitimately long phases (evaluation, checkpoint upload)."""
        return self._watchdog

    @property
    def gpu(self) -> bool:
        return self._status.gpu

    cuda = gpu

    @property
    def fp16(self) -> Optional[str]:
        return self._status.fp16

    @property
    def is_amp(self) -> bool:
        return self._status.is_amp

    @property
    def is_distributed(self) -> bool:
        return self._status.is_distributed

    @property
    def is_ddp(self) -> bool:
        return self._status.distributed == DistributedOptions.ddp.value

    @property
    def oss(self) -> bool:
        return self._status.oss

    @property
    def sharded(self) -> bool:
        return self._status.sharded

    @property
    def fully_sharded(self) -> bool:
        return self._status.fully_sharded

    @property
    def amp_config(self) -> AMPConfig:
        return self._amp_cfg

    @property
    def ddp_config(self) -> DDPConfig:
        return self._ddp_cfg

    @property
    def oss_config(self) -> FairscaleOSSConfig:
        return self._oss_cfg

    @property
    def sddp_config(self) -> FairscaleSDDPConfig:
        return self._sddp_cfg

    @property
    def fsdp_config(self) -> FairscaleFSDPConfig:
        return self._fsdp_cfg

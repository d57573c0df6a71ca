"""Typed configuration objects, option enums and the deferred optimizer spec.

Parity targets (reference call sites, all external-library surface):
  * ``AMPConfig(init_scale=2.**14)``                      Stoke-DDP.py:182-184
  * ``DDPConfig(local_rank=..., convert_to_sync_batch_norm=True)``  Stoke-DDP.py:190-193
  * ``FairscaleOSSConfig(broadcast_fp16=True)``           Stoke-DDP.py:197-199
  * ``ClipGradNormConfig(max_norm=..., norm_type=2.0)``   Stoke-DDP.py:253
  * ``StokeOptimizer(optimizer=AdamW, optimizer_kwargs={...})``  Stoke-DDP.py:226-235
  * ``DistributedOptions.ddp.value`` / ``FP16Options.amp.value``  Stoke-DDP.py:247-248
  * imported-but-unused ``DeepspeedConfig`` / ``DeepspeedZeROConfig``  Stoke-DDP.py:18

Configs are plain dataclasses.  The facade receives them as an untyped list
(Stoke-DDP.py:252) and dispatches on class, so every config class here is a
distinct type with no inheritance between the siblings.
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass
from enum import Enum
from typing import Any, Dict, List, Optional, Type, TypedDict

import torch


class DistributedOptions(Enum):
    """Distributed back-ends selectable on the facade (Stoke-DDP.py:248).

    Only ``ddp`` is implemented natively.  ``horovod`` and ``deepspeed`` are kept
    for API parity and raise ``NotImplementedError`` at facade construction.
    """

    ddp = "ddp"
    horovod = "horovod"
    deepspeed = "deepspeed"


class FP16Options(Enum):
    """Mixed-precision selectors (Stoke-DDP.py:247).

    ``amp`` is fp16 autocast + dynamic loss scaling.  ``bf16`` is an addition for
    B200 (no scaler needed).  The Apex and DeepSpeed members exist for parity and
    raise ``NotImplementedError``.
    """

    amp = "amp"
    bf16 = "bf16"
    apex_O1 = "apex_O1"
    apex_O2 = "apex_O2"
    deepspeed = "deepspeed"


class StokeOptimizer(TypedDict):
    """Deferred optimizer spec: the class plus its kwargs (Stoke-DDP.py:226-235).

    It is a *spec* so the facade can build the optimizer after the model has been
    moved, converted and (for OSS) partitioned.
    """

    optimizer: Type[torch.optim.Optimizer]
    optimizer_kwargs: Dict[str, Any]


@dataclass
class AMPConfig:
    """Dynamic loss-scaler settings (Stoke-DDP.py:182-184)."""

    backoff_factor: float = 0.5
    enabled: bool = True
    growth_factor: float = 2.0
    growth_interval: int = 2000
    init_scale: float = 2.0 ** 16


@dataclass
class DDPConfig:
    """Replicated data-parallel settings (Stoke-DDP.py:190-193)."""

    local_rank: Optional[int] = None
    auto_mpi_discovery: bool = False
    convert_to_sync_batch_norm: bool = False
    backend: str = "nccl"
    broadcast_buffers: bool = True
    bucket_cap_mb: int = 25
    find_unused_parameters: bool = False
    gradient_as_bucket_view: bool = True
    init_method: str = "env://"
    no_sync: bool = True
    # B200 additions
    fused_allreduce: bool = True  # use the peer-memory all-reduce kernel when available
    comm_dtype: Optional[torch.dtype] = None  # None: reduce in grad dtype; fp16/bf16 compresses the wire


@dataclass
class FairscaleOSSConfig:
    """Optimizer-state-sharding settings (Stoke-DDP.py:197-199)."""

    broadcast_fp16: bool = False


@dataclass
class FairscaleSDDPConfig:
    """Gradient-sharding (reduce-to-owner) settings."""

    auto_refresh_trainable: bool = True
    broadcast_buffers: bool = True
    reduce_buffer_size: int = 2 ** 23
    reduce_fp16: bool = False
    sync_models_at_startup: bool = True


@dataclass
class FairscaleFSDPConfig:
    """Full parameter sharding settings."""

    bucket_cap_mb: int = 25
    buffer_dtype: Optional[torch.dtype] = None
    clear_autocast_cache: bool = False
    compute_dtype: Optional[torch.dtype] = None
    flatten_parameters: bool = True
    force_input_to_fp32: bool = False
    fp32_reduce_scatter: bool = False
    gradient_predivide_factor: Optional[float] = None
    gradient_postdivide_factor: Optional[float] = None
    move_grads_to_cpu: Optional[bool] = None
    move_params_to_cpu: bool = False
    no_broadcast_optim_state: Optional[bool] = False
    reshard_after_forward: bool = True
    verbose: bool = False
    mixed_precision: bool = False
    min_num_params: int = int(1e6)  # auto-wrap threshold for child units


@dataclass
class ClipGradConfig:
    """Element-wise gradient value clipping."""

    clip_value: float = 1.0


@dataclass
class ClipGradNormConfig:
    """Global gradient-norm clipping (Stoke-DDP.py:253)."""

    max_norm: float = 1.0
    norm_type: float = 2.0


@dataclass
class ApexConfig:
    """Parity placeholder: Apex back-end is not implemented."""

    cast_model_outputs: Optional[torch.dtype] = None
    convert_to_sync_batch_norm: bool = False
    max_loss_scale: float = 2.0 ** 24
    min_loss_scale: Optional[float] = None
    scaler_per_loss: bool = False
    verbosity: int = 0


@dataclass
class HorovodConfig:
    """Parity placeholder: Horovod back-end is not implemented."""

    compression: bool = False
    convert_to_sync_batch_norm: bool = False
    gradient_predivide_factor: float = 1.0
    op: str = "Average"


@dataclass
class DeepspeedZeROConfig:
    """Parity placeholder (imported, unused, at Stoke-DDP.py:18)."""

    allgather_bucket_size: int = int(5e8)
    allgather_partitions: bool = True
    contiguous_gradients: bool = False
    overlap_comm: bool = False
    reduce_bucket_size: int = int(5e8)
    reduce_scatter: bool = True
    stage: int = 0


@dataclass
class DeepspeedConfig:
    """Parity placeholder (imported, unused, at Stoke-DDP.py:18)."""

    dist_backend: str = "nccl"
    fp32_allreduce: bool = False
    gradient_predivide_factor: float = 1.0
    prescale_gradients: bool = False
    zero_optimization: Optional[DeepspeedZeROConfig] = None
    zero_allow_untested_optimizer: bool = False


@dataclass
class B200Config:
    """Knobs that exist only in this framework (no reference counterpart).

    fused_kernels     use the sm_100a C-ABI kernel library for optimizer / scaler /
                      loss / norm when the tensors live on a CUDA device.
    peer_collectives  use symmetric-memory (peer / NVLS multicast) fused collectives
                      instead of NCCL for the hot paths.
    strict            a GPU run must really be on the native path: a missing kernel library or a
                      failed peer-memory attachment raises instead of falling back to stock ops /
                      NCCL (benchmarks and CI set this, so a fallback can never print a number).
    cuda_graph        capture the steady-state micro-step in a CUDA graph.
    channels_last     run convolutional models in NHWC.
    loss_sync_every   how often ``detach_and_sync_loss`` drains to the host
                      (1 keeps the reference's per-batch ``.item()`` semantics).
    loss_sync_lag     1: ``detach_and_sync_loss`` returns the previous call's value, so the host never waits
                      for the step it has just enqueued (0: wait, the reference's semantics).
    watchdog_timeout_s  spin limit of the in-kernel cross-rank barriers (a stuck peer raises a
                      status word instead of hanging the GPU).
    step_watchdog_s   host-side hang guard: if no optimizer step finishes for this many seconds
                      every thread's stack is dumped and, with ``step_watchdog_abort``, the process
                      exits non-zero so the launcher tears the job down (0 = off).
    """

    fused_kernels: bool = True
    peer_collectives: bool = True
    strict: bool = False
    cuda_graph: bool = False
    channels_last: bool = True
    loss_sync_every: int = 1
    loss_sync_lag: int = 0
    watchdog_timeout_s: float = 120.0
    step_watchdog_s: float = 0.0
    step_watchdog_abort: bool = False


ALL_CONFIG_TYPES = (
    AMPConfig,
    DDPConfig,
    FairscaleOSSConfig,
    FairscaleSDDPConfig,
    FairscaleFSDPConfig,
    ClipGradConfig,
    ClipGradNormConfig,
    ApexConfig,
    HorovodConfig,
    DeepspeedConfig,
    DeepspeedZeROConfig,
    B200Config,
)


def configs_to_dict(configs: Optional[List[Any]]) -> Dict[type, Any]:
    """Dispatch an untyped config list by class (Stoke-DDP.py:252)."""
    out: Dict[type, Any] = {}
    for cfg in configs or []:
        if not isinstance(cfg, ALL_CONFIG_TYPES):
            raise TypeError(
                f"Unknown config object {type(cfg).__name__}; expected one of "
                f"{[t.__name__ for t in ALL_CONFIG_TYPES]}"
            )
        if type(cfg) in out:
            raise ValueError(f"Duplicate config of type {type(cfg).__name__}")
        out[type(cfg)] = cfg
    return out


def asdict(cfg: Any) -> Dict[str, Any]:
    d = dataclasses.asdict(cfg)
    return {k: (str(v) if isinstance(v, torch.dtype) else v) for k, v in d.items()}

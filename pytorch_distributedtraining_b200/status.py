"""Flag-combination validation and the resolved run status.

Parity target: the checks Stoke runs on its constructor flags for the
call at Stoke-DDP.py:240-254: gradient sharding needs optimizer sharding; the
fairscale-style engines need ``distributed='ddp'``; full sharding excludes the other
two; AMP needs a GPU; Apex / DeepSpeed / Horovod are named for parity but rejected.
"""
from __future__ import annotations

from dataclasses import dataclass
from enum import Enum
from typing import Any, Dict, Optional, Union

from .configs import DistributedOptions, FP16Options


def _enum_value(x: Any, enum_cls) -> Optional[str]:
    """Accept an enum member, its ``.value`` string (what the reference passes,
    Stoke-DDP.py:247-248) or ``None``."""
    if x is None:
        return None
    if isinstance(x, enum_cls):
        return x.value
    if isinstance(x, Enum):
        return str(x.value)
    if isinstance(x, str):
        valid = {e.value for e in enum_cls}
        if x not in valid:
            raise ValueError(f"{x!r} is not a valid {enum_cls.__name__}; choose from {sorted(valid)}")
        return x
    raise TypeError(f"expected {enum_cls.__name__}, its value string, or None; got {type(x).__name__}")


@dataclass
class RunStatus:
    """What the facade resolved from its flags; saved in checkpoints as ``stoke_status``."""

    batch_size_per_device: int
    grad_accum_steps: int
    gpu: bool
    fp16: Optional[str]
    distributed: Optional[str]
    oss: bool
    sharded: bool
    fully_sharded: bool
    world_size: int = 1
    grad_clip: Optional[Dict[str, Any]] = None

    @property
    def is_amp(self) -> bool:
        return self.fp16 == FP16Options.amp.value

    @property
    def is_bf16(self) -> bool:
        return self.fp16 == FP16Options.bf16.value

    @property
    def is_distributed(self) -> bool:
        return self.distributed is not None

    @property
    def is_fairscale(self) -> bool:
        return self.oss or self.sharded or self.fully_sharded

    @property
    def effective_batch_size(self) -> int:
        return self.batch_size_per_device * self.grad_accum_steps * self.world_size

    def to_dict(self) -> Dict[str, Any]:
        return {
            "batch_size_per_device": self.batch_size_per_device,
            "grad_accum_steps": self.grad_accum_steps,
            "gpu": self.gpu,
            "fp16": self.fp16,
            "distributed": self.distributed,
            "fairscale_oss": self.oss,
            "fairscale_sddp": self.sharded,
            "fairscale_fsdp": self.fully_sharded,
            "world_size": self.world_size,
            "effective_batch_size": self.effective_batch_size,
            "grad_clip": self.grad_clip,
        }


def resolve_status(
    batch_size_per_device: int,
    grad_accum_steps: Optional[int],
    gpu: bool,
    fp16: Union[None, str, FP16Options],
    distributed: Union[None, str, DistributedOptions],
    fairscale_oss: bool,
    fairscale_sddp: bool,
    fairscale_fsdp: bool,
) -> RunStatus:
    fp16_v = _enum_value(fp16, FP16Options)
    dist_v = _enum_value(distributed, DistributedOptions)
    accum = 1 if grad_accum_steps is None else int(grad_accum_steps)
    if accum < 1:
        raise ValueError("grad_accum_steps must be >= 1")
    if int(batch_size_per_device) < 1:
        raise ValueError("batch_size_per_device must be >= 1")

    if dist_v in (DistributedOptions.horovod.value, DistributedOptions.deepspeed.value):
        raise NotImplementedError(
            f"distributed={dist_v!r} is kept for API parity only; this framework implements 'ddp' "
            f"(replicated, or sharded via fairscale_oss / fairscale_sddp / fairscale_fsdp)")
    if fp16_v in (FP16Options.apex_O1.value, FP16Options.apex_O2.value, FP16Options.deepspeed.value):
        raise NotImplementedError(f"fp16={fp16_v!r} is kept for API parity only; use 'amp' (fp16) or 'bf16'")
    if fp16_v == FP16Options.amp.value and not gpu:
        raise ValueError("fp16='amp' requires gpu=True (fp16 autocast with loss scaling is a CUDA path)")
    if fairscale_sddp and not fairscale_oss:
        raise ValueError("fairscale_sddp=True requires fairscale_oss=True: gradients are reduced to the rank "
                         "that owns the optimizer shard")
    if (fairscale_oss or fairscale_sddp or fairscale_fsdp) and dist_v != DistributedOptions.ddp.value:
        raise ValueError("fairscale_oss / fairscale_sddp / fairscale_fsdp require distributed='ddp'")
    if fairscale_fsdp and (fairscale_oss or fairscale_sddp):
        raise ValueError("fairscale_fsdp=True cannot be combined with fairscale_oss / fairscale_sddp: full "
                         "sharding already shards optimizer state and gradients")
    return RunStatus(
        batch_size_per_device=int(batch_size_per_device),
        grad_accum_steps=accum,
        gpu=bool(gpu),
        fp16=fp16_v,
        distributed=dist_v,
        oss=bool(fairscale_oss),
        sharded=bool(fairscale_sddp),
        fully_sharded=bool(fairscale_fsdp),
    )

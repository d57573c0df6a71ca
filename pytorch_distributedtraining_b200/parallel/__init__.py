"""Sharded data-parallel engines (SURVEY §1 layer L3)."""
from .comm import (barrier, destroy_process_group, find_free_port, get_rank, get_world_size,
                   init_process_group)
from .ddp import DistributedDataParallel, bf16_compress_hook, fp16_compress_hook
from .fsdp import FullyShardedDataParallel, auto_wrap, default_auto_wrap_policy
from .grad_scaler import GradScaler, ShardedGradScaler
from .oss import OSS
from .sharded_ddp import ShardedDataParallel
from .sync_bn import SyncBatchNorm, convert_sync_batchnorm

__all__ = [
    "OSS", "ShardedDataParallel", "DistributedDataParallel", "FullyShardedDataParallel", "auto_wrap",
    "default_auto_wrap_policy", "GradScaler", "ShardedGradScaler", "SyncBatchNorm", "convert_sync_batchnorm",
    "init_process_group", "destroy_process_group", "find_free_port", "get_rank", "get_world_size", "barrier",
    "fp16_compress_hook", "bf16_compress_hook",
]

"""B200-native sharded data-parallel training framework.

Public surface (what the two reference scripts import, under one roof):

    from pytorch_distributedtraining_b200 import (
        Stoke, StokeOptimizer, AMPConfig, DDPConfig, FairscaleOSSConfig, FairscaleSDDPConfig,
        FairscaleFSDPConfig, ClipGradConfig, ClipGradNormConfig, DistributedOptions, FP16Options,
        OSS, ShardedDataParallel, FullyShardedDataParallel, DistributedDataParallel)

See DESIGN.md for the architecture and INTEGRATION.md for the C ABI of the kernels.
"""
__version__ = "0.1.0"

from . import parallel  # noqa: F401
from .configs import (AMPConfig, ApexConfig, B200Config, ClipGradConfig, ClipGradNormConfig, DDPConfig,  # noqa: F401
                      DeepspeedConfig, DeepspeedZeROConfig, DistributedOptions, FairscaleFSDPConfig,
                      FairscaleOSSConfig, FairscaleSDDPConfig, FP16Options, HorovodConfig, StokeOptimizer)
from .data import (BucketedDistributedSampler, CustomDataset, DistributedSampler, PairedImageDataset,  # noqa: F401
                   StokeDataLoader, SyntheticImageDataset, SyntheticPairedDataset)
from .facade import Stoke  # noqa: F401
from .launch import find_free_port, spawn_training  # noqa: F401
from .parallel import (OSS, DistributedDataParallel, FullyShardedDataParallel, GradScaler,  # noqa: F401
                       ShardedDataParallel, ShardedGradScaler, SyncBatchNorm, convert_sync_batchnorm)

ShardedDDP = ShardedDataParallel
FSDP = FullyShardedDataParallel
DDP = DistributedDataParallel

"""Drop-in ``stoke`` namespace: put ``compat/`` on PYTHONPATH and the reference's
``from stoke import ...`` lines (Stoke-DDP.py:18-25) resolve to this framework."""
from pytorch_distributedtraining_b200 import (AMPConfig, ApexConfig, ClipGradConfig, ClipGradNormConfig, DDPConfig,  # noqa: F401
                                              DeepspeedConfig, DeepspeedZeROConfig, DistributedOptions,
                                              FairscaleFSDPConfig, FairscaleOSSConfig, FairscaleSDDPConfig,
                                              FP16Options, HorovodConfig, Stoke, StokeOptimizer)
from pytorch_distributedtraining_b200.data import BucketedDistributedSampler  # noqa: F401

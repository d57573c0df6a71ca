from pytorch_distributedtraining_b200.parallel import FullyShardedDataParallel, ShardedDataParallel  # noqa: F401

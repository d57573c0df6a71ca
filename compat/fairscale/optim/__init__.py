from pytorch_distributedtraining_b200.parallel import OSS, GradScaler, ShardedGradScaler  # noqa: F401

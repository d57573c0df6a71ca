from pytorch_distributedtraining_b200.parallel.grad_scaler import GradScaler, ShardedGradScaler  # noqa: F401

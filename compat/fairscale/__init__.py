"""Drop-in ``fairscale`` namespace (``fairscale.optim.oss.OSS``,
``fairscale.nn.data_parallel.ShardedDataParallel`` / ``FullyShardedDataParallel``;
Fairscale-DDP.py:6-7)."""

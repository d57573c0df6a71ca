from pytorch_distributedtraining_b200.parallel.oss import OSS  # noqa: F401

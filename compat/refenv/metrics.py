"""``metrics.mae`` / ``metrics.psnr`` (Stoke-DDP.py:120-121)."""
from pytorch_distributedtraining_b200.utils.metrics import accuracy, mae, mse, psnr  # noqa: F401

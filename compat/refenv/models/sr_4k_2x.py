"""``from models.sr_4k_2x import Net`` (Fairscale-DDP.py:13)."""
from pytorch_distributedtraining_b200.models.sr_net import Net  # noqa: F401

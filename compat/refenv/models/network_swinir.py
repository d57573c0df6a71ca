"""``from models.network_swinir import SwinIR`` (Stoke-DDP.py:33)."""
from pytorch_distributedtraining_b200.models.swinir import SwinIR  # noqa: F401

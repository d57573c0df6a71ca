"""``from PyTorchPercept import feat_loss`` (Stoke-DDP.py:35,224)."""
from pytorch_distributedtraining_b200.models.perceptual import PerceptualLoss, feat_loss  # noqa: F401

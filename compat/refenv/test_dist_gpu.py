"""``from test_dist_gpu import *`` (Fairscale-DDP.py:18): the script takes ``os`` and ``find_free_port`` from it."""
import os  # noqa: F401

import _torch_shim  # noqa: F401
from pytorch_distributedtraining_b200.launch import find_free_port  # noqa: F401

__all__ = ["os", "find_free_port"]

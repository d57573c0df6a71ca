"""B200-native sharded data-parallel training framework."""
__version__ = "0.1.0"
from . import parallel  # noqa: F401
from .configs import *  # noqa: F401,F403

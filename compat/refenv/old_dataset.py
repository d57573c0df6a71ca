"""Stand-in for the reference's private ``old_dataset.CustomDataset(input_path, target_path)`` (Fairscale-DDP.py:37,
Stoke-DDP.py:264): paired low / high resolution patches.  Real directories are read when they exist; otherwise
(the hard-coded share paths of the scripts) deterministic synthetic pairs of the same structure are served."""
import os

import _torch_shim  # noqa: F401
from pytorch_distributedtraining_b200.data import PairedImageDataset, SyntheticPairedDataset


class CustomDataset:
    def __new__(cls, input_path, target_path, *a, **k):
        if os.path.isdir(str(input_path)) and os.path.isdir(str(target_path)):
            return PairedImageDataset(input_path, target_path)
        n = int(os.environ.get("REFENV_SAMPLES", "64"))
        p = int(os.environ.get("REFENV_PATCH", "64"))
        return SyntheticPairedDataset(n, lr_size=p, scale=2)

"""Offline stand-in for ``wandb`` (Stoke-DDP.py:42-58,316-325,339): same calls, JSON lines on disk, no network.
The reference logs from every rank; here only rank 0 writes."""
import os

import _torch_shim  # noqa: F401
from pytorch_distributedtraining_b200.utils.logging import RunLogger

_run = None
config = {}


def login(*a, **k):
    return True


def init(project=None, config=None, reinit=False, **k):
    global _run
    if _run is None or reinit:
        _run = RunLogger(project=project or "refenv", config=dict(config or {}), rank=int(os.environ.get("RANK", "0") or 0),
                         use_wandb=False)
    if config:
        globals()["config"] = dict(config)
    return _run


def log(data, **k):
    if _run is not None:
        _run.log({key: (float(v) if hasattr(v, "item") else v) for key, v in data.items()})


def finish(*a, **k):
    global _run
    if _run is not None and hasattr(_run, "finish"):
        _run.finish()
    _run = None

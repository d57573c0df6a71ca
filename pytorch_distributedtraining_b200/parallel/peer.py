"""Attachment point for the peer-memory (NVLink / NVSwitch) fused collectives.

Filled in by ``parallel.symm`` + ``ops.collectives`` when a symmetric-memory arena can
be created for the process group; until then every hook reports "not attached" and
the engines use their ``torch.distributed`` collectives (the baseline).
"""
from __future__ import annotations

from typing import Any, Optional

import torch

_STATE: dict = {"engine": None}


def attach(engine: Any, optim: Any, status: Any, **cfg: Any) -> Optional[Any]:
    from . import symm
    return symm.attach(engine, optim, status, **cfg)


def small_allreduce(buf: torch.Tensor, group: Any) -> bool:
    """Sum ``buf`` (a few KB of fp32) across ranks in place with the one-shot peer kernel.
    Returns False when no arena is attached for ``group``."""
    eng = _STATE.get("engine")
    if eng is None:
        return False
    return eng.small_allreduce(buf, group)


def set_engine(engine: Any) -> None:
    _STATE["engine"] = engine

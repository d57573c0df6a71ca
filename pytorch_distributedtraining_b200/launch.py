"""Process bootstrap: the two launch styles the reference uses.

  * ``mp.spawn(train, args=(WORLD_SIZE, EPOCHS), nprocs=WORLD_SIZE, join=True)`` with
    ``MASTER_ADDR=127.0.0.1`` and ``MASTER_PORT=find_free_port()``
    (Fairscale-DDP.py:112-133) → :func:`spawn_training`;
  * ``python -m torch.distributed.launch --nproc_per_node=N script.py``
    (Stoke-DDP.py:1-2), which exports ``LOCAL_RANK`` / ``RANK`` / ``WORLD_SIZE`` →
    :func:`env_rank_info` reads them (``torchrun`` sets the same variables).

Fixes relative to the reference: ``os`` is actually imported, ``rank`` is
derived rather than computed-and-dropped, and the world size is a parameter instead of a
hard-coded 4.
"""
from __future__ import annotations

import os
from typing import Any, Callable, Tuple

import torch
import torch.multiprocessing as mp

from .parallel.comm import find_free_port

__all__ = ["find_free_port", "spawn_training", "env_rank_info"]


def env_rank_info() -> Tuple[int, int, int]:
    """(rank, local_rank, world_size) from the launcher's environment; a missing or
    literal ``"None"`` LOCAL_RANK (the reference writes that string, Stoke-DDP.py:153)
    means single-process."""
    def geti(name: str, default: int) -> int:
        v = os.environ.get(name)
        return default if v in (None, "", "None") else int(v)
    return geti("RANK", 0), geti("LOCAL_RANK", 0), geti("WORLD_SIZE", 1)


def _child(rank: int, fn: Callable[..., Any], world_size: int, args: tuple) -> None:
    os.environ["RANK"] = str(rank)
    os.environ["LOCAL_RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(world_size)
    fn(rank, world_size, *args)


def spawn_training(fn: Callable[..., Any], world_size: int, args: tuple = (), master_addr: str = "127.0.0.1",
                   master_port: str = "") -> None:
    """Run ``fn(rank, world_size, *args)`` in ``world_size`` processes on this host.

    ``fn`` is responsible for ``init_process_group`` (as in the reference's ``train``),
    or can simply construct a ``Stoke`` object, which does it.
    """
    os.environ["MASTER_ADDR"] = master_addr
    os.environ["MASTER_PORT"] = master_port or find_free_port()
    mp.spawn(_child, args=(fn, world_size, args), nprocs=world_size, join=True)

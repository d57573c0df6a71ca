"""Multi-process test harness: N ranks on one host over gloo/loopback.

Pattern from the reference's own launcher (``mp.spawn`` + ``MASTER_ADDR=127.0.0.1`` +
a free port, Fairscale-DDP.py:122-133), wrapped so a test is one function call.
"""
from __future__ import annotations

import os
from typing import Any, Callable

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ..parallel.comm import find_free_port


def _entry(rank: int, world_size: int, port: str, backend: str, fn: Callable, args: tuple) -> None:
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = port
    os.environ["RANK"] = str(rank)
    os.environ["LOCAL_RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(world_size)
    torch.manual_seed(1234)
    torch.set_num_threads(1)
    if backend == "nccl":
        torch.cuda.set_device(rank)
    dist.init_process_group(backend=backend, init_method="env://", rank=rank, world_size=world_size)
    try:
        fn(rank, world_size, *args)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def spawn(fn: Callable[..., Any], world_size: int = 2, args: tuple = (), backend: str = "gloo") -> None:
    """Run ``fn(rank, world_size, *args)`` on ``world_size`` processes; raises if any rank fails."""
    port = find_free_port()
    mp.spawn(_entry, args=(world_size, port, backend, fn, args), nprocs=world_size, join=True)

"""Process-group plumbing shared by every engine.

Reference call sites: ``dist.init_process_group(backend='gloo', init_method="env://",
rank=rank, world_size=world_size)`` (Fairscale-DDP.py:27), ``destroy_process_group``
(Fairscale-DDP.py:109); NCCL + ``env://`` implied by the Stoke DDP config
(Stoke-DDP.py:190-193).

Design: one process per GPU, ``torch.distributed`` for rendezvous and cold paths
(initial parameter broadcast, object gathers for checkpoints, CPU/gloo plumbing).
Hot-path collectives on CUDA are routed to the peer-memory kernels in
``ops.collectives`` when a symmetric-memory arena has been attached to the group
(see ``parallel.symm``); otherwise they fall back to the backend collective, which
is the measured baseline.
"""
from __future__ import annotations

import datetime
import os
import socket
from contextlib import closing
from typing import Any, List, Optional

import torch
import torch.distributed as dist


def find_free_port() -> str:
    """Free TCP port on loopback as a string (``find_free_port()`` at Fairscale-DDP.py:123)."""
    with closing(socket.socket(socket.AF_INET, socket.SOCK_STREAM)) as s:
        s.bind(("127.0.0.1", 0))
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        return str(s.getsockname()[1])


def is_initialized() -> bool:
    return dist.is_available() and dist.is_initialized()


def get_rank(group: Optional[Any] = None) -> int:
    return dist.get_rank(group) if is_initialized() else 0


def get_world_size(group: Optional[Any] = None) -> int:
    return dist.get_world_size(group) if is_initialized() else 1


def get_global_rank(group: Optional[Any], group_rank: int) -> int:
    """Translate a rank inside ``group`` to the global rank collectives want."""
    if group is None or group is dist.group.WORLD or not is_initialized():
        return group_rank
    return dist.get_global_rank(group, group_rank)


def init_process_group(
    backend: Optional[str] = None,
    init_method: str = "env://",
    rank: Optional[int] = None,
    world_size: Optional[int] = None,
    timeout_s: float = 1800.0,
    device_id: Optional[torch.device] = None,
) -> None:
    """Idempotent ``dist.init_process_group`` with a finite watchdog timeout.

    ``backend=None`` picks NCCL when CUDA is present, else gloo.  A hung rank makes the
    collective raise after ``timeout_s`` instead of hanging the job forever (the
    reference has no timeout).
    """
    if is_initialized():
        return
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if rank is None:
        rank = int(os.environ.get("RANK", "0"))
    if world_size is None:
        world_size = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    kwargs = {}
    if device_id is not None and backend == "nccl":
        kwargs["device_id"] = device_id
    dist.init_process_group(
        backend=backend,
        init_method=init_method,
        rank=rank,
        world_size=world_size,
        timeout=datetime.timedelta(seconds=timeout_s),
        **kwargs,
    )


def destroy_process_group() -> None:
    if is_initialized():
        dist.destroy_process_group()


def barrier(group: Optional[Any] = None) -> None:
    if is_initialized() and get_world_size(group) > 1:
        if dist.get_backend(group) == "nccl":
            dist.barrier(group=group, device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier(group=group)


def broadcast_object(obj: Any, src: int = 0, group: Optional[Any] = None) -> Any:
    if not is_initialized() or get_world_size(group) == 1:
        return obj
    box = [obj]
    dist.broadcast_object_list(box, src=get_global_rank(group, src), group=group)
    return box[0]


def all_gather_object(obj: Any, group: Optional[Any] = None) -> List[Any]:
    if not is_initialized() or get_world_size(group) == 1:
        return [obj]
    out: List[Any] = [None] * get_world_size(group)
    dist.all_gather_object(out, obj, group=group)
    return out


@torch.no_grad()
def broadcast_coalesced(tensors: List[torch.Tensor], src: int = 0, group: Optional[Any] = None,
                        buffer_elems: int = 1 << 24) -> None:
    """Broadcast many tensors from ``src`` with few collectives (cold path, K11).

    Tensors are packed per (dtype, device) into flat staging buffers of at most
    ``buffer_elems`` elements, broadcast, and scattered back.
    """
    if not is_initialized() or get_world_size(group) == 1 or not tensors:
        return
    gsrc = get_global_rank(group, src)
    by_kind = {}
    for t in tensors:
        by_kind.setdefault((t.dtype, t.device), []).append(t)
    for (_dtype, _dev), ts in by_kind.items():
        chunk: List[torch.Tensor] = []
        n = 0
        def flush():
            nonlocal chunk, n
            if not chunk:
                return
            if len(chunk) == 1 and chunk[0].is_contiguous():
                dist.broadcast(chunk[0], src=gsrc, group=group)
            else:
                flat = torch.cat([c.reshape(-1) for c in chunk])
                dist.broadcast(flat, src=gsrc, group=group)
                off = 0
                for c in chunk:
                    c.copy_(flat[off:off + c.numel()].view_as(c))
                    off += c.numel()
            chunk, n = [], 0
        for t in ts:
            if n + t.numel() > buffer_elems and chunk:
                flush()
            chunk.append(t)
            n += t.numel()
        flush()

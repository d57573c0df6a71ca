"""ctypes bindings for the peer-memory collectives of ``csrc/collectives.cu``.

A :class:`PeerContext` wraps the C ``b200dp_peer_ctx``.  Buffers are described by
*pointer tables*: entry ``p`` is rank ``p``'s buffer as mapped into this process.  The
same bindings drive the real multi-GPU path (tables from torch symmetric memory, see
``parallel.symm``) and the single-GPU emulation used by the tests (tables of local
tensors, ``emulate=True``).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch

from ._lib import check, require_library
from .fused import DT, AdamHParams

MAX_RANKS = 8
# measurement only (bench.py "exposed_comm_ms"): when True every collective launch below is skipped, so the same
# step can be timed with its communication removed.  Results are garbage while it is set.
MUTE = False
VoidPtrTable = ctypes.c_void_p * MAX_RANKS


class PeerCtxStruct(ctypes.Structure):
    _fields_ = [("rank", ctypes.c_int), ("world", ctypes.c_int), ("emulate", ctypes.c_int),
                ("max_blocks", ctypes.c_int),
                ("flags", ctypes.c_void_p * MAX_RANKS), ("epochs", ctypes.c_void_p * MAX_RANKS),
                ("status", ctypes.c_void_p * MAX_RANKS), ("timeout_ns", ctypes.c_uint64),
                ("skew_ns", ctypes.c_uint64)]


def table(ptrs: Optional[Sequence[Optional[int]]]) -> Optional[VoidPtrTable]:
    """Pointer table from ints / tensors / None (missing entries are NULL)."""
    if ptrs is None:
        return None
    t = VoidPtrTable()
    for i, p in enumerate(ptrs):
        if p is None:
            t[i] = None
        elif isinstance(p, torch.Tensor):
            t[i] = p.data_ptr()
        else:
            t[i] = int(p)
    return t


def local_table(rank: int, x: Optional[torch.Tensor]) -> Optional[VoidPtrTable]:
    """Table with only this rank's entry filled (rank-local objects)."""
    if x is None:
        return None
    t = VoidPtrTable()
    t[rank] = x.data_ptr()
    return t


class PeerContext:
    """Flags, epochs and status words for one process group + launch helpers."""

    def __init__(self, rank: int, world: int, flag_ptrs: Sequence[int], epochs: torch.Tensor, status: torch.Tensor,
                 max_blocks: int, emulate: bool = False, timeout_s: float = 30.0,
                 epoch_ptrs: Optional[Sequence[int]] = None, status_ptrs: Optional[Sequence[int]] = None):
        self.rank, self.world, self.emulate, self.max_blocks = rank, world, emulate, max_blocks
        self.epochs, self.status = epochs, status  # keep alive
        c = PeerCtxStruct()
        c.rank, c.world, c.emulate, c.max_blocks = rank, world, int(emulate), max_blocks
        for p in range(world):
            c.flags[p] = int(flag_ptrs[p])
        if emulate:
            for p in range(world):
                c.epochs[p] = int(epoch_ptrs[p])
                c.status[p] = int(status_ptrs[p])
        else:
            c.epochs[rank] = epochs.data_ptr()
            c.status[rank] = status.data_ptr()
        c.timeout_ns = int(timeout_s * 1e9)
        self.c = c
        self.lib = require_library()

    def _s(self, stream: Optional[torch.cuda.Stream]) -> ctypes.c_void_p:
        return ctypes.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)

    def timed_out(self) -> bool:
        """Host read of the watchdog word (synchronises; call at checkpoints, not per step)."""
        return bool(self.status.max().item() != 0)

    # -- collectives -----------------------------------------------------------
    def barrier(self, blocks: int = 1, rounds: int = 1, stream=None) -> None:
        check(self.lib.b200dp_peer_barrier(ctypes.byref(self.c), blocks, rounds, self._s(stream)), "peer_barrier")

    def allreduce(self, bufs: VoidPtrTable, mc: int, n: int, dtype: torch.dtype, pre_mult: float = 1.0,
                  mult_dev: Optional[VoidPtrTable] = None, found_inf: Optional[VoidPtrTable] = None, algo: int = 1,
                  blocks: int = 32, stream=None) -> None:
        if MUTE:
            return
        check(self.lib.b200dp_peer_allreduce(ctypes.byref(self.c), bufs, ctypes.c_void_p(mc or None), ctypes.c_int64(n),
                                             DT[dtype], ctypes.c_float(pre_mult), mult_dev, found_inf, algo, blocks,
                                             self._s(stream)), "peer_allreduce")

    def reduce_scatter(self, in_bufs: VoidPtrTable, mc: int, n_per_rank: int, dtype: torch.dtype, outs: VoidPtrTable,
                       pre_mult: float = 1.0, mult_dev: Optional[VoidPtrTable] = None,
                       found_inf: Optional[VoidPtrTable] = None, accumulate: bool = False, blocks: int = 32,
                       stream=None) -> None:
        if MUTE:
            return
        check(self.lib.b200dp_peer_reduce_scatter(ctypes.byref(self.c), in_bufs, ctypes.c_void_p(mc or None),
                                                  ctypes.c_int64(n_per_rank), DT[dtype], outs, ctypes.c_float(pre_mult),
                                                  mult_dev, found_inf, int(accumulate), blocks, self._s(stream)),
              "peer_reduce_scatter")

    def all_gather(self, outs: VoidPtrTable, mc: int, shards: VoidPtrTable, n_per_rank: int, shard_dtype: torch.dtype,
                   out_dtype: torch.dtype, blocks: int = 32, stream=None) -> None:
        if MUTE:
            return
        check(self.lib.b200dp_peer_all_gather(ctypes.byref(self.c), outs, ctypes.c_void_p(mc or None), shards,
                                              ctypes.c_int64(n_per_rank), DT[shard_dtype], DT[out_dtype], blocks,
                                              self._s(stream)), "peer_all_gather")

    def zero_adam(self, grad_bufs: Optional[VoidPtrTable], grad_mc: int, param_bufs: VoidPtrTable, param_mc: int,
                  seg: Sequence[int], wire_dtype: torch.dtype, master: VoidPtrTable, m: VoidPtrTable, v: VoidPtrTable,
                  hp: AdamHParams, step: VoidPtrTable, pre_mult: float = 1.0, mult_dev: Optional[VoidPtrTable] = None,
                  found_inf: Optional[VoidPtrTable] = None, prereduced: bool = False,
                  local_grad: Optional[VoidPtrTable] = None, blocks: int = 32, stream=None) -> None:
        if MUTE:
            return
        seg_arr = (ctypes.c_int64 * (MAX_RANKS + 1))(*list(seg) + [seg[-1]] * (MAX_RANKS + 1 - len(seg)))
        check(self.lib.b200dp_peer_zero_adam(ctypes.byref(self.c), grad_bufs, ctypes.c_void_p(grad_mc or None), param_bufs,
                                             ctypes.c_void_p(param_mc or None), seg_arr, DT[wire_dtype], master, m, v, hp,
                                             step, ctypes.c_float(pre_mult), mult_dev, found_inf, int(prereduced),
                                             local_grad, blocks, self._s(stream)), "peer_zero_adam")

    def reduce_to_owner(self, grad_bufs: VoidPtrTable, seg: Sequence[int], local_out: VoidPtrTable,
                        pre_mult: float, mult_dev: Optional[VoidPtrTable], norm_out: VoidPtrTable,
                        partials: VoidPtrTable, counters: VoidPtrTable, blocks: int = 32, stream=None) -> None:
        if MUTE:
            return
        seg_arr = (ctypes.c_int64 * (MAX_RANKS + 1))(*list(seg) + [seg[-1]] * (MAX_RANKS + 1 - len(seg)))
        check(self.lib.b200dp_peer_reduce_to_owner(ctypes.byref(self.c), grad_bufs, seg_arr, local_out,
                                                   ctypes.c_float(pre_mult), mult_dev, norm_out, partials, counters,
                                                   blocks, self._s(stream)), "peer_reduce_to_owner")

    def broadcast(self, bufs: VoidPtrTable, mc: int, nbytes: int, root: int = 0, blocks: int = 8, stream=None) -> None:
        if MUTE:
            return
        check(self.lib.b200dp_peer_broadcast(ctypes.byref(self.c), bufs, ctypes.c_void_p(mc or None), ctypes.c_int64(nbytes),
                                             root, blocks, self._s(stream)), "peer_broadcast")

    def small_allreduce(self, staging: VoidPtrTable, io: VoidPtrTable, n: int, slot_floats: int,
                        call_counter: VoidPtrTable, stream=None) -> None:
        if MUTE:
            return
        check(self.lib.b200dp_peer_small_allreduce(ctypes.byref(self.c), staging, io, n, slot_floats, call_counter,
                                                   self._s(stream)), "peer_small_allreduce")


class EmulatedPeers:
    """All ranks on ONE GPU: per-rank flag / epoch / status arrays and a context with
    ``emulate=True``.  Every kernel launch is cooperative with ``grid.y = world``."""

    def __init__(self, world: int, max_blocks: int = 16, device: Optional[torch.device] = None, timeout_s: float = 10.0):
        device = device or torch.device("cuda", torch.cuda.current_device())
        self.world, self.device = world, device
        # rows [0, max_blocks): barrier flags; rows [max_blocks, 2*max_blocks): non-finite notices of the NVLS all-reduce
        self.flags = [torch.zeros(2 * max_blocks * MAX_RANKS, dtype=torch.int32, device=device) for _ in range(world)]
        self.epochs = [torch.zeros(max_blocks, dtype=torch.int32, device=device) for _ in range(world)]
        self.status = [torch.zeros(1, dtype=torch.int32, device=device) for _ in range(world)]
        self.ctx = PeerContext(0, world, [f.data_ptr() for f in self.flags], self.epochs[0], self.status[0], max_blocks,
                               emulate=True, timeout_s=timeout_s, epoch_ptrs=[e.data_ptr() for e in self.epochs],
                               status_ptrs=[s.data_ptr() for s in self.status])

    def timed_out(self) -> bool:
        return any(int(s.item()) != 0 for s in self.status)

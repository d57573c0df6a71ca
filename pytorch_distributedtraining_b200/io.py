"""Checkpoint save / load for the facade.

Parity target: ``path, tag = stoke_model.save(path='checkpoint/', name='model_...')``
(Stoke-DDP.py:142-145).  One ``torch.save`` dictionary with the keys
``backward_step, grad_accum_step, optimizer_step, stoke_status, model_state_dict,
optimizer_state_dict, scaler_state_dict, extras``; tag
``stoke-{name}-backward-step-{n}``; sharded optimizer state is consolidated to rank 0
before writing; with full sharding every rank participates in the gather; rank 0
writes.  The model and optimizer entries are *plain torch layouts* (un-prefixed module
keys; optimizer state keyed by global parameter index), so a checkpoint loads into an
un-wrapped ``nn.Module`` + stock ``torch.optim`` instance and into any world size.

The reference never resumes (``--start-epoch`` is dead, Stoke-DDP.py:161); ``load``
is provided because the facade surface has it and SURVEY §5.4 asks for it.
"""
from __future__ import annotations

import os
from typing import Any, Dict, Optional, Tuple

import torch
import torch.distributed as dist

from .parallel import comm
from .parallel.fsdp import FullyShardedDataParallel
from .parallel.oss import OSS


def make_tag(name: str, backward_step: int) -> str:
    return f"stoke-{name}-backward-step-{backward_step}"


def _strip_prefix(sd: Dict[str, Any], prefix: str = "module.") -> Dict[str, Any]:
    if sd and all(k.startswith(prefix) for k in sd):
        return {k[len(prefix):]: v for k, v in sd.items()}
    return sd


def _fsdp_optim_state(stoke) -> Dict[str, Any]:
    """Name-keyed, un-flattened optimizer state (world-size independent)."""
    return stoke._model.gather_full_optim_state_dict(stoke._optimizer)


def _load_fsdp_optim_state(stoke, full: Dict[str, Any]) -> None:
    model: FullyShardedDataParallel = stoke._model
    optim = stoke._optimizer
    for g, saved in zip(optim.param_groups, full.get("param_groups", [])):
        for k, v in saved.items():
            if k != "params":
                g[k] = v
    for uname, unit in model.named_modules():
        if not isinstance(unit, FullyShardedDataParallel) or unit.flat_param is None:
            continue
        prefix = (uname + "." if uname else "").replace("module.", "")
        layout = unit.param_layout()
        keys = set()
        for fq, *_ in layout:
            keys |= set(full["state"].get(prefix + fq, {}).keys())
        st: Dict[str, Any] = {}
        for key in keys:
            sample = full["state"][prefix + layout[0][0]][key]
            if isinstance(sample, torch.Tensor) and sample.dim() > 0:
                flat = torch.zeros(unit._padded, dtype=sample.dtype, device=unit.flat_param.device)
                for fq, _shape, off, n in layout:
                    flat[off: off + n].copy_(full["state"][prefix + fq][key].reshape(-1))
                st[key] = flat[unit.rank * unit._shard_numel:(unit.rank + 1) * unit._shard_numel].clone()
            else:
                st[key] = sample.clone() if isinstance(sample, torch.Tensor) else sample
        if st:
            optim.state[unit.flat_param] = st


def save(stoke, path: str, name: str = "", extension: str = "pt", create_directory: bool = True,
         extras: Optional[dict] = None) -> Tuple[str, str]:
    tag = make_tag(name, stoke._backward_steps)
    full_path = os.path.join(path, f"{tag}.{extension}")
    rank = stoke.rank
    model, optim = stoke._model, stoke._optimizer

    # collectives first: every rank must take part
    if isinstance(model, FullyShardedDataParallel):
        model_sd = model.state_dict()
        optim_sd = _fsdp_optim_state(stoke)
        optim_layout = "fsdp_named"
    else:
        if hasattr(model, "_flush_lazy_buffers"):
            model._flush_lazy_buffers()      # BatchNorm statistics whose per-forward broadcast was elided (collective)
        model_sd = _strip_prefix(stoke._unwrapped.state_dict())
        if isinstance(optim, OSS):
            optim.consolidate_state_dict(recipient_rank=0)
            optim_sd = optim.state_dict() if rank == 0 else None
        else:
            optim_sd = optim.state_dict()
        optim_layout = "torch"

    if rank == 0:
        if create_directory:
            os.makedirs(path, exist_ok=True)
        payload = {
            "backward_step": stoke._backward_steps,
            "grad_accum_step": stoke._grad_accum_counter,
            "optimizer_step": stoke._optimizer_steps,
            "stoke_status": stoke.status,
            "model_state_dict": {k: v.detach().cpu() for k, v in model_sd.items()},
            "optimizer_state_dict": optim_sd,
            "optimizer_state_layout": optim_layout,
            "scaler_state_dict": stoke._scaler.state_dict() if stoke._scaler is not None else None,
            "extras": extras,
        }
        tmp = full_path + ".tmp"
        torch.save(payload, tmp)
        os.replace(tmp, full_path)  # atomic: a crash never leaves a half-written checkpoint
    comm.barrier()
    return path, tag


def load(stoke, path: str, tag: str, strict: bool = True, extension: str = "pt") -> Optional[dict]:
    full_path = os.path.join(path, f"{tag}.{extension}")
    ckpt = torch.load(full_path, map_location="cpu", weights_only=False)
    model, optim = stoke._model, stoke._optimizer
    model_sd = _strip_prefix(ckpt["model_state_dict"])
    if isinstance(model, FullyShardedDataParallel):
        model.load_state_dict(model_sd, strict=strict)
        if ckpt.get("optimizer_state_layout") == "fsdp_named" and ckpt["optimizer_state_dict"] is not None:
            _load_fsdp_optim_state(stoke, ckpt["optimizer_state_dict"])
    else:
        stoke._unwrapped.load_state_dict(model_sd, strict=strict)
        if ckpt["optimizer_state_dict"] is not None:
            optim.load_state_dict(ckpt["optimizer_state_dict"])
    if stoke._scaler is not None and ckpt.get("scaler_state_dict"):
        stoke._scaler.load_state_dict(ckpt["scaler_state_dict"])
    stoke._backward_steps = int(ckpt["backward_step"])
    stoke._grad_accum_counter = int(ckpt["grad_accum_step"])
    stoke._optimizer_steps = int(ckpt["optimizer_step"])
    comm.barrier()
    return ckpt.get("extras")

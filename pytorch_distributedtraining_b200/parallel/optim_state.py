"""Consolidation of optimizer state that lives on FSDP flat shards.

Parity target: fairscale ``FSDP.gather_full_optim_state_dict``:
turn per-rank state tensors shaped like the flat shard into per-original-parameter
tensors so a checkpoint is independent of the world size.
"""
from __future__ import annotations

from typing import Any, Dict

import torch
import torch.distributed as dist


def gather_full_optim_state(root, optim: torch.optim.Optimizer) -> Dict[str, Any]:
    from .fsdp import FullyShardedDataParallel, _all_gather_into

    out_state: Dict[str, Dict[str, Any]] = {}
    for uname, unit in root.named_modules():
        if not isinstance(unit, FullyShardedDataParallel) or unit.flat_param is None:
            continue
        st = optim.state.get(unit.flat_param, {})
        prefix = uname + "." if uname else ""
        prefix = prefix.replace("module.", "")
        for key, val in st.items():
            if isinstance(val, torch.Tensor) and val.numel() == unit._shard_numel:
                full = torch.empty(unit._padded, dtype=val.dtype, device=val.device)
                _all_gather_into(full, val.reshape(-1).contiguous(), unit.process_group, unit.world_size)
                for fq, shape, off, n in unit.param_layout():
                    out_state.setdefault(prefix + fq, {})[key] = full[off: off + n].view(shape).clone()
            else:
                for fq, _shape, _off, _n in unit.param_layout():
                    out_state.setdefault(prefix + fq, {})[key] = val.clone() if isinstance(val, torch.Tensor) else val
    groups = [{k: v for k, v in g.items() if k != "params"} for g in optim.param_groups]
    return {"state": out_state, "param_groups": groups}

"""Run logging: Weights & Biases when importable and reachable, else a JSONL file.

Parity target: ``wandb.login`` / ``wandb.init(project, config, reinit=True)`` with a
retry loop / ``wandb.log({...})`` / ``wandb.finish`` (Stoke-DDP.py:42-58,308-325,339),
``tqdm`` epoch bar (Stoke-DDP.py:330) and the rank-filtered prints.

Deliberate fixes (SURVEY §7 "reference bugs not to replicate"): one ``init`` per run
instead of one per log call; rank-0-only logging; bounded retries instead of a bare
``except`` spinning forever; works offline (there is no network on the build box).
"""
from __future__ import annotations

import json
import os
import time
from typing import Any, Dict, Optional


class RunLogger:
    def __init__(self, project: str = "b200-sharded-dp", config: Optional[Dict[str, Any]] = None, rank: int = 0,
                 log_dir: str = "runs", use_wandb: Optional[bool] = None, retries: int = 3, retry_sleep_s: float = 2.0):
        self.rank = rank
        self.project = project
        self.config = dict(config or {})
        self._wandb = None
        self._fh = None
        self._t0 = time.time()
        if rank != 0:
            return
        want = use_wandb if use_wandb is not None else os.environ.get("WANDB_MODE", "") not in ("disabled", "offline")
        if want:
            try:
                import wandb  # type: ignore
                for attempt in range(retries):
                    try:
                        wandb.init(project=project, config=self.config, reinit=True)
                        self._wandb = wandb
                        break
                    except Exception as e:  # network is optional
                        print(f"[logger] wandb.init failed ({e!r}); retry {attempt + 1}/{retries}")
                        time.sleep(retry_sleep_s)
            except ImportError:
                self._wandb = None
        if self._wandb is None:
            os.makedirs(log_dir, exist_ok=True)
            path = os.path.join(log_dir, f"{project}-{int(self._t0)}.jsonl")
            self._fh = open(path, "a")
            self._fh.write(json.dumps({"event": "init", "project": project, "config": self.config}) + "\n")
            self.path = path

    def log(self, payload: Dict[str, Any], step: Optional[int] = None) -> None:
        if self.rank != 0:
            return
        clean = {k: (float(v) if hasattr(v, "item") else v) for k, v in payload.items()}
        if self._wandb is not None:
            self._wandb.log(clean, step=step)
        elif self._fh is not None:
            clean["_t"] = round(time.time() - self._t0, 3)
            if step is not None:
                clean["_step"] = step
            self._fh.write(json.dumps(clean) + "\n")
            self._fh.flush()

    def finish(self) -> None:
        if self._wandb is not None:
            self._wandb.finish()
        if self._fh is not None:
            self._fh.close()
            self._fh = None


def train_log(logger: RunLogger, loss: Any, example_ct: int, epoch: int) -> None:
    """Stoke-DDP.py:47-52."""
    logger.log({"epoch": epoch, "train_loss": loss})
    if logger.rank == 0:
        print(f"Loss after {str(example_ct).zfill(5)} examples: {float(loss):.3f}")


def val_log(logger: RunLogger, loss: Any, avg_mae: Any, avg_psnr: Any, example_ct: int, epoch: int) -> None:
    """Stoke-DDP.py:54-59."""
    logger.log({"epoch": epoch, "val_loss": loss, "PSNR": avg_psnr, "MAE": avg_mae})
    if logger.rank == 0:
        print(f"-----VALIDATION Loss after {str(example_ct).zfill(5)} examples: {float(loss):.3f}--------")

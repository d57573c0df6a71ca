"""Device-side timing, clock sampling and NVTX ranges (the reference has none of this,
SURVEY §5.1; BASELINE's metric is device-timed, max over ranks).
"""
from __future__ import annotations

import contextlib
import statistics
import subprocess
import threading
import time
from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist


class CudaStepTimer:
    """CUDA-event timer on the current stream; ``elapsed_ms`` synchronises."""

    def __init__(self) -> None:
        self._start = torch.cuda.Event(enable_timing=True)
        self._stop = torch.cuda.Event(enable_timing=True)

    def start(self) -> None:
        self._start.record()

    def stop(self) -> None:
        self._stop.record()

    def elapsed_ms(self) -> float:
        self._stop.synchronize()
        return self._start.elapsed_time(self._stop)


def max_over_ranks(value: float, device: Optional[torch.device] = None) -> float:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([value], dtype=torch.float64, device=device or (
            torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else "cpu"))
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return value


@contextlib.contextmanager
def nvtx_range(name: str):
    if torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples ``nvidia-smi`` clocks and throttle reasons in a background thread during
    a timed region (B200_PROFILING.md "clocks DURING the timed region")."""

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.gpu_index, self.period_s = gpu_index, period_s
        self.rows: List[List[str]] = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None

    def _run(self) -> None:
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits",
                                      "-i", str(self.gpu_index)], capture_output=True, text=True, timeout=5).stdout
                for line in out.strip().splitlines():
                    self.rows.append([c.strip() for c in line.split(",")])
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def start(self) -> "ClockSampler":
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict[str, Any]:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=10)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, val in zip(names, r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        return {
            "sm_mhz": statistics.median(sm) if sm else None,
            "sm_max_mhz": max(mx) if mx else None,
            "reasons": sorted(reasons),
            "samples": len(sm),
        }


def exposed_comm_ms(engine: Any, fwd_bwd: Any, iters: int = 10, warmup: int = 3) -> Dict[str, float]:
    """Communication that is NOT hidden behind compute, in ms per forward+backward.

    ``engine`` is a data-parallel wrapper with ``no_sync()`` (DDP, ShardedDDP, FSDP), ``fwd_bwd()`` runs one
    forward + backward through it (and should clear the gradients itself if it wants to).  The same work is
    timed with gradient synchronisation on and, under ``no_sync()``, off; the difference is what the collectives
    add to the critical path.  Device-timed with CUDA events on a GPU, wall-clock on the CPU; max over ranks.
    """
    cuda = torch.cuda.is_available() and any(p.is_cuda for p in engine.parameters())
    dev = next((p.device for p in engine.parameters()), torch.device("cpu"))

    def timed(n: int) -> float:
        if cuda:
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                fwd_bwd()
            e1.record()
            e1.synchronize()
            return e0.elapsed_time(e1) / n
        t0 = time.perf_counter()
        for _ in range(n):
            fwd_bwd()
        return (time.perf_counter() - t0) * 1e3 / n

    for _ in range(warmup):
        fwd_bwd()
    sync_ms = timed(iters)
    with engine.no_sync():
        for _ in range(max(1, warmup // 2)):
            fwd_bwd()
        nosync_ms = timed(iters)
    fwd_bwd()          # one synchronising pass so replicas leave with reduced gradients again
    sync_ms = max_over_ranks(sync_ms, dev if cuda else None)
    nosync_ms = max_over_ranks(nosync_ms, dev if cuda else None)
    return {"sync_ms": sync_ms, "no_sync_ms": nosync_ms, "exposed_comm_ms": max(sync_ms - nosync_ms, 0.0)}

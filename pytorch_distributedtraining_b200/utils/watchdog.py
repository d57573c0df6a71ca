"""Host-side step watchdog (the reference has no hang detection — a stuck rank hangs the job).

The in-kernel barriers already give up after ``B200Config.watchdog_timeout_s`` and raise a status word; this
is the matching guard on the host: a daemon thread that expects a ``beat()`` at least every ``timeout_s``
seconds (the facade beats once per optimizer step).  When the beats stop it dumps every Python thread's stack
(``faulthandler``), calls an optional callback, and — if ``abort`` — ends the process with a non-zero exit
code so the launcher (``torchrun`` / ``launch.spawn_training``) tears the job down instead of waiting forever.
"""
from __future__ import annotations

import faulthandler
import os
import sys
import threading
import time
import traceback
from typing import Callable, Optional

__all__ = ["StepWatchdog"]


class StepWatchdog:
    def __init__(self, timeout_s: float, on_timeout: Optional[Callable[[float], None]] = None, abort: bool = False,
                 exit_code: int = 3, name: str = "b200dp-step-watchdog", stream=None):
        if timeout_s <= 0:
            raise ValueError("timeout_s must be positive")
        self.timeout_s, self.on_timeout, self.abort, self.exit_code = float(timeout_s), on_timeout, abort, exit_code
        self._stream = stream if stream is not None else sys.stderr
        self._last = time.monotonic()
        self._beats = 0
        self._fired = 0
        self._paused = False
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, name=name, daemon=True)
        self._thread.start()

    # ------------------------------------------------------------------ producer side
    def beat(self) -> None:
        """Progress mark: call once per step (cheap: one clock read)."""
        self._last = time.monotonic()
        self._beats += 1

    def pause(self) -> None:
        """Suspend the check (checkpointing, evaluation, anything legitimately long)."""
        self._paused = True

    def resume(self) -> None:
        self._last = time.monotonic()
        self._paused = False

    @property
    def beats(self) -> int:
        return self._beats

    @property
    def fired(self) -> int:
        return self._fired

    def close(self) -> None:
        self._stop.set()
        if self._thread.is_alive() and threading.current_thread() is not self._thread:
            self._thread.join(timeout=1.0)

    def __enter__(self) -> "StepWatchdog":
        return self

    def __exit__(self, *exc) -> None:
        self.close()

    # ------------------------------------------------------------------ watcher thread
    def _dump_stacks(self) -> None:
        try:
            self._stream.fileno()
            faulthandler.dump_traceback(file=self._stream, all_threads=True)      # async-signal-safe C path
        except Exception:
            # in-memory / wrapped streams have no descriptor: format the frames in Python instead
            names = {t.ident: t.name for t in threading.enumerate()}
            for ident, frame in sys._current_frames().items():
                print(f"Thread {names.get(ident, '?')} ({ident}):", file=self._stream)
                print("".join(traceback.format_stack(frame)), file=self._stream, flush=True)

    def _run(self) -> None:
        poll = min(1.0, self.timeout_s / 4)
        while not self._stop.wait(poll):
            if self._paused:
                continue
            idle = time.monotonic() - self._last
            if idle < self.timeout_s:
                continue
            self._fired += 1
            rank = os.environ.get("RANK", "0")
            try:
                print(f"[b200dp watchdog] rank {rank}: no training step finished for {idle:.1f} s "
                      f"(limit {self.timeout_s:.1f} s, {self._beats} steps so far); thread stacks follow",
                      file=self._stream, flush=True)
                self._dump_stacks()
            except Exception:          # a closed stream must not take the watcher down
                pass
            if self.on_timeout is not None:
                try:
                    self.on_timeout(idle)
                except Exception:
                    pass
            if self.abort:
                os._exit(self.exit_code)
            self._last = time.monotonic()      # report again after another full period

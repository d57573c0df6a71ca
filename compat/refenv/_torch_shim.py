"""The reference scripts target an older torch: restore what they pass and current torch rejects.
  * ``ReduceLROnPlateau(..., verbose=True)`` (Stoke-DDP.py:302-306): the keyword was removed.
  * ``REFENV_MAX_WORKERS``: the scripts hard-code 16 / 8 DataLoader workers per rank (Fairscale-DDP.py:59-64,
    Stoke-DDP.py:297); on a small box cap them (0 = load in the main process)."""
import functools
import os

import torch
import torch.utils.data

_done = False


def apply() -> None:
    global _done
    if _done:
        return
    _done = True
    sched = torch.optim.lr_scheduler.ReduceLROnPlateau
    orig_init = sched.__init__

    @functools.wraps(orig_init)
    def init(self, *a, verbose=None, **k):
        return orig_init(self, *a, **k)

    sched.__init__ = init
    cap = os.environ.get("REFENV_MAX_WORKERS")
    if cap is not None:
        cap = int(cap)
        dl = torch.utils.data.DataLoader
        dl_init = dl.__init__

        @functools.wraps(dl_init)
        def dinit(self, *a, **k):
            if k.get("num_workers", 0) > cap:
                k["num_workers"] = cap
            if k.get("num_workers", 0) == 0:
                for key in ("multiprocessing_context", "prefetch_factor", "persistent_workers"):
                    k.pop(key, None)
            return dl_init(self, *a, **k)

        dl.__init__ = dinit


apply()

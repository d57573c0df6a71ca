"""tcgen05 GEMM front end (filled in with csrc/gemm_sm100.cu)."""
from __future__ import annotations


def available() -> bool:
    return False

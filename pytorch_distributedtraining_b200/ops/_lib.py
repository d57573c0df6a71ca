"""Build and load the sm_100a kernel library.

The kernels are plain CUDA behind a C ABI (``include/b200dp.h``), compiled by nvcc into
``ops/libb200dp.so`` *in-tree* (so the built library travels with a repository snapshot)
and loaded with ctypes.  No torch headers are involved: a full rebuild takes seconds
and the ``.so`` has no Python or ATen dependency.

On a machine with a GPU the library is mandatory: ``require_library`` raises if it is
missing rather than letting a run silently fall back to stock ops.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List, Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
_CSRC = os.path.join(_ROOT, "csrc")
_INCLUDE = os.path.join(_ROOT, "include")
_BUILD = os.path.join(_ROOT, "build", "b200dp")
_SO = os.path.join(_HERE, "libb200dp.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "--use_fast_math", "-Xcompiler", "-fPIC", "-Xcompiler", "-O3", "-Xptxas", "-v",
    "--expt-relaxed-constexpr",
]

_lib: Optional[ctypes.CDLL] = None
_load_error: Optional[str] = None


def lib_path() -> str:
    return _SO


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (looked at $NVCC, /usr/local/cuda/bin/nvcc, PATH)")


def _sources() -> List[str]:
    return sorted(os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith(".cu"))


def _stale(target: str, deps: List[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libb200dp.so``.  Works without a
    GPU (nvcc cross-compiles).  Incremental: only stale objects are rebuilt."""
    os.makedirs(_BUILD, exist_ok=True)
    nvcc = _nvcc()
    headers = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".h"))]
    headers += [os.path.join(_INCLUDE, f) for f in os.listdir(_INCLUDE)]
    srcs = _sources()
    objs = [os.path.join(_BUILD, os.path.basename(s)[:-3] + ".o") for s in srcs]

    def compile_one(pair) -> Optional[str]:
        src, obj = pair
        if not force and not _stale(obj, [src] + headers):
            return None
        cmd = [nvcc, *NVCC_FLAGS, "-I", _INCLUDE, "-I", _CSRC, "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = os.path.join(_BUILD, os.path.basename(src) + ".log")
        with open(log, "w") as fh:
            fh.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs) or 1)) as ex:
        rebuilt = [o for o in ex.map(compile_one, zip(srcs, objs)) if o]
    if rebuilt or force or _stale(_SO, objs):
        cmd = [nvcc, "-shared", "-o", _SO + ".tmp", *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        os.replace(_SO + ".tmp", _SO)
    global _lib, _load_error
    _lib, _load_error = None, None
    return _SO


def _try_load() -> Optional[ctypes.CDLL]:
    global _lib, _load_error
    if _lib is not None:
        return _lib
    if _load_error is not None:
        return None
    if not os.path.exists(_SO):
        _load_error = f"{_SO} does not exist; run `python -c 'import __graft_entry__ as g; g.build()'`"
        return None
    try:
        lib = ctypes.CDLL(_SO, mode=ctypes.RTLD_GLOBAL)
        lib.b200dp_last_error.restype = ctypes.c_char_p
        lib.b200dp_version.restype = ctypes.c_int
        for name in ("b200dp_sumsq_scratch_floats", "b200dp_bn_scratch_floats", "b200dp_ln_scratch_floats",
                     "b200dp_column_sums_scratch_floats"):
            getattr(lib, name).restype = ctypes.c_int64
        _lib = lib
    except OSError as e:  # e.g. no libcuda on a CPU-only box
        _load_error = f"cannot load {_SO}: {e}"
        return None
    return _lib


def library_available() -> bool:
    """True when the kernel library is built, loadable, and a CUDA device is present."""
    import torch
    return torch.cuda.is_available() and _try_load() is not None


def require_library() -> ctypes.CDLL:
    lib = _try_load()
    if lib is None:
        raise RuntimeError(f"b200dp kernel library unavailable: {_load_error}")
    return lib


LAUNCHES = 0  # kernels launched through this library in this process (bench.py reports it)


def check(rc: int, what: str = "", launches: int = 1) -> None:
    global LAUNCHES
    LAUNCHES += launches
    if rc != 0:
        lib = require_library()
        msg = lib.b200dp_last_error().decode(errors="replace")
        raise RuntimeError(f"b200dp {what} failed (code {rc}): {msg}")


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))

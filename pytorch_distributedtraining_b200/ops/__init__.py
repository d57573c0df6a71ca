"""Hand-written sm_100a kernels behind a C ABI (``include/b200dp.h``), loaded with
ctypes from the in-tree ``libb200dp.so``; see ``_lib.py`` for build and loading."""
from ._lib import library_available, require_library, build_library, lib_path  # noqa: F401

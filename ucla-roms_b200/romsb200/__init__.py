"""romsb200 -- thin ctypes host binding of libromsb200.so (the C ABI in include/roms_b200.h).

The product is the CUDA library; this module only mirrors its surface for Python hosts (tests,
bench.py, __graft_entry__.py).  It never computes on the CPU: if the shared library is missing
the import of `Roms` users fails loudly (RuntimeError), there is no fallback.
"""
from .capi import RomsCfg, FIELDS, FID, CApi, Roms, lib_path, load_library  # noqa: F401
from . import cases  # noqa: F401

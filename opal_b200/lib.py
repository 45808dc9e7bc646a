"""Loads the product library libopal_b200.so (CUDA, sm_100a) and binds its C ABI.

There is deliberately no fallback: if the shared object is missing the import of `load()`
raises, and on a machine without a CUDA device `opal_create` returns OPAL_ECUDA.
"""
from __future__ import annotations

import ctypes
import os
import pathlib

from .abi import OpalABI

_HERE = pathlib.Path(__file__).resolve().parent
LIB_PATH = _HERE / "csrc" / "libopal_b200.so"
_abi = None


def load() -> OpalABI:
    global _abi
    if _abi is None:
        path = pathlib.Path(os.environ.get("OPAL_B200_LIB", LIB_PATH))   # development: alternative builds
        if not path.exists():
            raise RuntimeError(f"{path} not built: run `make -C {_HERE / 'csrc'}` "
                               "(or python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
        _abi = OpalABI(ctypes.CDLL(os.fspath(path), mode=ctypes.RTLD_GLOBAL), "opal_")
    return _abi

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
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} not built: run `make -C {_HERE / 'csrc'}` "
                               "(or python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
        _abi = OpalABI(ctypes.CDLL(os.fspath(LIB_PATH), mode=ctypes.RTLD_GLOBAL), "opal_")
    return _abi

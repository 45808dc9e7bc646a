import ctypes
import os
import pathlib
import subprocess
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _build_oracle():
    so = ROOT / "oracle" / "libopal_oracle.so"
    src = ROOT / "oracle" / "opal_oracle.cpp"
    if not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
        subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    return so


@pytest.fixture(scope="session")
def oracle_cdll():
    return ctypes.CDLL(os.fspath(_build_oracle()))


@pytest.fixture(scope="session")
def oracle(oracle_cdll):
    """The CPU oracle bound through the same ABI class as the product (prefix opal_ref_)."""
    from opal_b200.abi import OpalABI
    return OpalABI(oracle_cdll, "opal_ref_")


@pytest.fixture(scope="session")
def gpu():
    """The product library on a CUDA device; fails loudly if it is not built."""
    from opal_b200 import lib
    return lib.load()

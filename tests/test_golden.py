"""Committed golden vectors (tests/golden/*.npz, written by tests/golden/make_golden.py from the CPU oracle
after it passed the reference's known-answer tests).

CPU: the oracle still reproduces them (1e-12: its deposit sums per OpenMP chunk, so the last bits may move
with the thread count, as in the reference itself, yee.rs:621-628; integer work bit-exact).
GPU: the CUDA path, through the C ABI, lands on the same vectors -- keys and counts bit-exact, floats within
north_star's 1e-10 -- without the oracle or /root/reference present at run time.
"""
import pathlib
import sys

import numpy as np
import pytest

import pic_cases as pc

GOLDEN = pathlib.Path(__file__).resolve().parent / "golden"
sys.path.insert(0, str(GOLDEN))
import make_golden as mg  # noqa: E402

CASES = sorted(mg.CASES)


def _check(name, abi, tol, qtol):
    want = np.load(GOLDEN / f"{name}.npz")
    got = mg.CASES[name](abi)
    assert sorted(got) == sorted(want.files)
    for k in want.files:
        if k.endswith("_cell") or k.endswith("_weight"):
            assert np.array_equal(got[k], want[k]), f"{name}: {k} differs (bit-exact contract)"
    for k in want.files:
        if k in ("a_survivors", "a_e_cell", "a_ph_cell", "a_ph_weight"):
            assert np.array_equal(got[k], want[k]), f"{name}: {k} differs (same photons absorbed by the same electrons)"
        elif k.endswith("_defined"):
            assert np.array_equal(got[k], want[k]), f"{name}: {k} differs (Some / None pattern)"
        elif k.startswith("q_"):          # pure functions spanning many decades: element-wise relative error
            nz = want[k] != 0.0
            assert np.array_equal(got[k][~nz], want[k][~nz])
            err = np.max(np.abs(got[k][nz] / want[k][nz] - 1.0))
            assert err <= qtol, f"{name}: {k} off by {err} element-wise"
        elif k == "fields":
            pc.compare_fields(got[k], want[k], tol=tol)
        elif not (k.endswith("_cell") or k.endswith("_weight")):
            err = pc.rel_inf(got[k], want[k])
            assert err <= tol, f"{name}: {k} off by {err}"


def test_fixtures_are_committed():
    for name in CASES:
        assert (GOLDEN / f"{name}.npz").is_file(), f"run tests/golden/make_golden.py ({name})"


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_golden(oracle, name):
    _check(name, oracle, 1.0e-12, 1.0e-14)


@pytest.mark.gpu
@pytest.mark.parametrize("name", [c for c in CASES if c not in mg.CPU_ONLY])
def test_cuda_matches_golden(gpu, name):
    _check(name, gpu, pc.FLOAT_TOL, 1.0e-12)     # the bound of tests/test_gpu_kat.py against the live oracle

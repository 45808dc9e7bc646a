"""Pins the CPU oracle against every known-answer test the reference holds for the hot path
(SURVEY 8c).  The oracle is only trusted as a parity checker because these pass."""
import ctypes as C
import math

import numpy as np
import pytest

import kat


def test_table_checksum(oracle_cdll):
    oracle_cdll.opal_ref_table_checksum.restype = C.c_uint64
    oracle_cdll.opal_ref_table_checksum_expected.restype = C.c_uint64
    assert oracle_cdll.opal_ref_table_checksum() == oracle_cdll.opal_ref_table_checksum_expected()


def test_laser_bc(oracle):
    em, target, ey = kat.laser_bc(oracle)
    assert abs(em - target) / target < 1.0e-3
    # printed (not asserted) by the reference: Ey/E0 at lambda/4 and 5 lambda/4
    assert ey[0] == pytest.approx(-math.cos(math.pi / 32.0) ** 2, abs=2e-2)
    assert ey[1] == pytest.approx(-math.cos(5.0 * math.pi / 32.0) ** 2, abs=2e-2)


def test_conducting_bc(oracle):
    em, target = kat.conducting_bc(oracle)
    assert abs(em - target) / target < 1.0e-3


def test_periodic_bc(oracle):
    e1, e2 = kat.periodic_bc(oracle)
    assert e1 > 0.0
    assert abs(e1 - e2) / e1 < 1.0e-6


def test_static_magnetic_field(oracle):
    p, target = kat.static_magnetic_field(oracle)
    assert (p["gamma"][0] - target) / target < 1.0e-6          # the reference's one-sided check
    assert abs(p["gamma"][0] - target) / target < 1.0e-6       # and two-sided
    # after one full period the momentum is back along +x
    assert p["ux"][0] == pytest.approx(1.0, abs=5e-3)


def test_static_electric_field(oracle):
    p, target = kat.static_electric_field(oracle)
    assert (p["gamma"][0] - target) / target < 1.0e-6
    assert abs(p["gamma"][0] - target) / target < 1.0e-6


def test_rate_kats(oracle):
    got, want = kat.rate_kats(oracle)
    assert np.all(np.abs(got - want) / want < 1.0e-3)


def _invert(cdll, y, table):
    cdll.opal_ref_pwmci_invert.argtypes = [C.c_double, C.c_void_p, C.c_int64, C.POINTER(C.c_double), C.POINTER(C.c_int)]
    t = np.ascontiguousarray(table, dtype=np.float64)
    x, n = C.c_double(), C.c_int()
    ok = cdll.opal_ref_pwmci_invert(y, t.ctypes.data, len(t), C.byref(x), C.byref(n))
    assert ok == 1
    return x.value, n.value


def test_pwmci_invert_x_sqd(oracle_cdll):       # qed/pwmci.rs:235-248
    table = [[i / 20.0, (i / 20.0) ** 2] for i in range(20)]
    x, _ = _invert(oracle_cdll, 0.73, table)
    assert abs(x - math.sqrt(0.73)) < 1.0e-4


def test_pwmci_invert_tanh(oracle_cdll):        # qed/pwmci.rs:250-264
    table = [[5.0 * i / 40.0, math.tanh(5.0 * i / 40.0)] for i in range(40)]
    x, _ = _invert(oracle_cdll, 0.22, table)
    assert abs(x - math.atanh(0.22)) < 1.0e-4


def test_pwmci_invert_shifted_tanh(oracle_cdll):  # qed/pwmci.rs:266-281
    table = [[5.0 * i / 20.0, 1.0 + math.tanh(5.0 * i / 20.0 - 2.0)] for i in range(20)]
    x, _ = _invert(oracle_cdll, 1.24, table)
    assert abs(x - (2.0 - math.atanh(1.0 - 1.24))) < 1.0e-4


def test_vec3_orthogonality(oracle_cdll):        # vec3.rs:152-158
    a = (C.c_double * 3)(1.0, 2.0, -3.0)          # reference draws a random vector
    out = (C.c_double * 3)()
    oracle_cdll.opal_ref_vec3_orthogonal(a, out)
    assert abs(sum(a[i] * out[i] for i in range(3))) < 1.0e-10
    assert abs(sum(v * v for v in out) - 1.0) < 1.0e-10


def test_vec3_rotation(oracle_cdll):             # vec3.rs:160-169: two pi/2 rotations take x to z
    oracle_cdll.opal_ref_vec3_rotate_around.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
    v = (C.c_double * 3)(1.0, 0.0, 0.0)
    zaxis = (C.c_double * 3)(0.0, 0.0, 1.0)
    xaxis = (C.c_double * 3)(1.0, 0.0, 0.0)
    o1, o2 = (C.c_double * 3)(), (C.c_double * 3)()
    oracle_cdll.opal_ref_vec3_rotate_around(v, zaxis, math.pi / 2.0, o1)       # x -> y
    oracle_cdll.opal_ref_vec3_rotate_around(o1, xaxis, math.pi / 2.0, o2)      # y -> z
    assert abs(o2[0]) < 1.0e-10 and abs(o2[1]) < 1.0e-10 and abs(o2[2] - 1.0) < 1.0e-10


def test_airy_kats(oracle):                      # qed/special_functions/airy.rs:77-112
    rel, ok = kat.airy_kats(oracle)
    assert max(rel) < 1.0e-12, rel
    assert list(ok) == [True, True, True, True, False, False, False, False]      # None: x = 200, x < 0, x >= 50, NaN


def test_airy_against_an_independent_evaluation(oracle):
    """the restated algorithm (series below 1, 40/16/4-node Gauss-Laguerre above) against scipy's Ai over [0, 50)"""
    from scipy.special import airy
    x = np.concatenate([np.linspace(0.0, 2.0, 801), np.linspace(2.0, 49.99, 1500)])
    ai, ok = oracle.airy(x)
    assert ok.all()
    assert np.max(np.abs(ai - airy(x)[0]) / airy(x)[0]) < 2.0e-12


def test_pair_cross_sections_branches(oracle):
    """photon_absorption.rs:17-35 / stimulated_emission.rs:17-37 against a direct numpy statement of the formulas"""
    from scipy.special import airy
    k4, p4, cg, ce = kat.pair_cases()
    sa, ss, oka, oks = oracle.pair_cross_sections(k4, p4, cg, ce)
    # same association as the reference: k.p cancels to ~1e-8 of k0 p0 for these pairs and Ai amplifies what is left
    kp = k4[:, 0] * p4[:, 0] - k4[:, 1] * p4[:, 1] - k4[:, 2] * p4[:, 2] - k4[:, 3] * p4[:, 3]
    r2 = (2.0 * np.pi * 2.817940e-15) ** 2
    seen = set()
    for sign, sig, ok in ((+1, sa, oka), (-1, ss, oks)):
        with np.errstate(all="ignore"):
            den = ce * (ce + sign * cg)
            z = np.abs(cg / den) ** (2.0 / 3.0)
            zbar = 2.0 * z * ce * kp / cg
            expect_ok = (ce > 0) & (cg > 0) & (zbar >= 0) & (zbar < 50.0)
            if sign < 0:
                expect_ok &= (cg < ce) & (k4[:, 0] < p4[:, 0])
            g = 0.5 + 0.25 * cg ** 2 / den
            val = r2 * ce * z * (4.0 * g * (2.0 * p4[:, 0] * kp / k4[:, 0]) - 1.0) * airy(zbar)[0] / (7.29735257e-3 * cg * k4[:, 0] * p4[:, 0])
        assert (ok == expect_ok).all()
        assert (sig[~ok] == 0.0).all()
        assert np.max(np.abs(sig[ok] - val[ok]) / np.abs(val[ok])) < 1.0e-11
        seen |= set(np.digitize(zbar[ok], [1.0, 2.0, 10.0]))
        assert ok.sum() > 500 and (~ok).sum() > 100
    assert seen == {0, 1, 2, 3}                   # every range of the Airy evaluation was exercised

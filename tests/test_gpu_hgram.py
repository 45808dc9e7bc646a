"""Device-side distribution functions (opal_hgram_minmax / opal_hgram_bin) against the CPU statement of
particle/hgram.rs on identical particles.  Bin INDICES are integer work: for the algebraic quantities (x, p, energy)
the min/max are bit-identical and every particle lands in the same bin, so the weighted counts agree to summation
round-off; for quantities that go through libm (hypot, acos, atan2, ln) a value on a bin edge may move by one bin."""
import math

import numpy as np
import pytest

from opal_b200.abi import ELECTRON, ION, PHOTON, FLAG_LEAN_STATE, W_ENERGY, W_WEIGHT, Slab
import pic_cases as pc

pytestmark = pytest.mark.gpu


def _grid(mn, mx, n):
    """number_of_bins + bin_size_and_volume of hgram.rs:130-164 (Automatic)"""
    nb, sz = [], []
    for a, b in zip(mn, mx):
        k = 1 if a == b else int(math.ceil(2.0 * np.cbrt(float(n))))
        nb.append(k)
        sz.append(0.0 if a == b else (b - a) / k)
    return nb, sz


def _compare(a, b, species, quantities, exact, log_scaled=None, weight=W_WEIGHT):
    mna, mxa, na = a.hgram_minmax(species, quantities, log_scaled)
    mnb, mxb, nb_ = b.hgram_minmax(species, quantities, log_scaled)
    assert na == nb_
    if exact:
        assert np.array_equal(mna, mnb) and np.array_equal(mxa, mxb), (quantities, mna, mnb)
    else:
        assert np.allclose(mna, mnb, rtol=1e-14, atol=0) and np.allclose(mxa, mxb, rtol=1e-14, atol=0)
    nbins, sz = _grid(mnb, mxb, nb_)
    if len(quantities) == 1 and sz[0] == 0.0:
        sz[0] = 1.0                                     # hgram.rs:209-211
    ca, ta = a.hgram_bin(species, quantities, mnb, sz, nbins, log_scaled, weight)
    cb, tb = b.hgram_bin(species, quantities, mnb, sz, nbins, log_scaled, weight)
    assert abs(ta - tb) <= 1e-12 * abs(tb)
    assert abs(ca.sum() - cb.sum()) <= 1e-12 * abs(cb.sum())
    if exact:
        assert np.allclose(ca, cb, rtol=1e-11, atol=1e-13 * np.abs(cb).max()), quantities
    else:
        moved = np.abs(ca - cb).sum() / np.abs(cb).sum()
        assert moved < 1e-3, (quantities, moved)
    # the last bin edge: the particle holding the maximum is dropped or aliased, exactly as hgram.rs:253 / 357
    return ca, cb


@pytest.mark.parametrize("flags", [0, FLAG_LEAN_STATE])
def test_electron_histograms_match_oracle(gpu, oracle, flags):
    nx, npc, dx = 96, 150, 500.0
    a = Slab(gpu, nx, -1.0e4, dx, flags=flags)
    b = Slab(oracle, nx, -1.0e4, dx)
    p = pc.thermal_particles(nx, npc, dx, 0.3, seed=3)
    p["weight"] = p["weight"] * (1.0 + 0.5 * np.random.default_rng(4).random(p["weight"].size))
    pc.upload(a, ELECTRON, p)
    pc.upload(b, ELECTRON, p)
    for q in (["x"], ["px"], ["py"], ["pz"], ["energy"], ["x", "px"], ["py", "pz"], ["x", "energy"]):
        _compare(a, b, ELECTRON, q, exact=True)
    _compare(a, b, ELECTRON, ["energy"], exact=True, weight=W_ENERGY)
    for q in (["p_perp"], ["theta"], ["phi"], ["x", "p_perp"], ["longitude", "latitude"]):
        _compare(a, b, ELECTRON, q, exact=False)
    _compare(a, b, ELECTRON, ["energy"], exact=False, log_scaled=[1])
    _compare(a, b, ELECTRON, ["energy", "energy"], exact=False, log_scaled=[1, 1], weight=W_ENERGY)
    if flags == 0:
        a.run_steps(3)
        b.run_steps(3)          # work, y, z, chi become non-trivial; states now differ by <= 1e-10
        for q in (["chi"], ["work"], ["r"]):
            mna, mxa, _ = a.hgram_minmax(ELECTRON, q)
            mnb, mxb, _ = b.hgram_minmax(ELECTRON, q)
            assert np.allclose(mna, mnb, rtol=1e-8, atol=1e-300) and np.allclose(mxa, mxb, rtol=1e-8, atol=1e-300), q


def test_ion_and_photon_histograms_match_oracle(gpu, oracle):
    nx, dx = 64, 1.0e-8
    a = Slab(gpu, nx, 0.0, dx)
    b = Slab(oracle, nx, 0.0, dx)
    p = pc.thermal_particles(nx, 80, dx, 0.01, seed=8, ne=1.0e24)
    rng = np.random.default_rng(12)
    n = 3000
    k = rng.standard_normal((n, 3)) * 50.0
    ph_cell, ph_x = np.sort(rng.integers(0, nx, n)), rng.random(n)
    for s_ in (a, b):
        s_.set_ion(6.0, 12.0)
        pc.upload(s_, ION, p)
        s_.upload(PHOTON, ph_cell, ph_x, k[:, 0], k[:, 1], k[:, 2], np.full(n, 2.5))
    for q in (["x"], ["px"], ["energy"], ["x", "px"]):
        _compare(a, b, ION, q, exact=True)
        _compare(a, b, PHOTON, q, exact=True)
    _compare(a, b, PHOTON, ["longitude", "latitude"], exact=False, weight=W_ENERGY)
    _compare(a, b, PHOTON, ["energy", "energy"], exact=False, log_scaled=[1, 1], weight=W_ENERGY)
    # empty species: (DBL_MAX, -DBL_MAX, 0) like hgram.rs:189
    c = Slab(gpu, nx, 0.0, dx)
    mn, mx, cnt = c.hgram_minmax(ELECTRON, ["x"])
    assert cnt == 0 and mn[0] == np.finfo(float).max and mx[0] == -np.finfo(float).max

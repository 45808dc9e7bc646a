"""The oracle's absorb (particle/interactions.rs:145-340) against the properties the reference's algorithm has: what a
pair loop without events leaves behind, exact momentum bookkeeping of absorption and stimulated emission, the filters,
and the reference's untouched last occupied cell."""
import numpy as np

from opal_b200.abi import ELECTRON, FLAG_NO_STIMULATED_EMISSION, PHOTON
import absorb_cases as ac


def _momentum(el, ph):
    """sum of weight x momentum / mc over both species, per component"""
    return np.array([np.sum(el["weight"] * el[k]) + np.sum(ph["weight"] * ph[k]) for k in ("ux", "uy", "uz")])


def test_pair_loop_without_events_only_spends_optical_depth(oracle):
    s = ac.build(oracle)
    el0, ph0, _, _ = ac.snapshot(s)
    n = ph0["cell"].size
    ac.set_depths(s, np.full(n, 50.0), np.full(n, 60.0))
    s.absorb(0.0)
    el1, ph1, ta, ts = ac.snapshot(s)
    for k in el0:
        assert np.array_equal(el0[k], el1[k]), k
    for k in ph0:
        assert np.array_equal(ph0[k], ph1[k], equal_nan=True), k
    has_e = np.isin(ph0["cell"], el0["cell"]) & (ph0["cell"] != el0["cell"].max())
    assert np.all(ta[~has_e] == 50.0) and np.all(ts[~has_e] == 60.0)          # no electrons there, or the last occupied cell
    assert np.all(ta[has_e] < 50.0)                                           # absorption is always possible for chi > 0
    assert np.any(ts[has_e] < 60.0) and np.any(ts[has_e] == 60.0)             # stimulated emission needs k0 < gamma, chi_g < chi_e
    assert np.all(ts >= 0.0) and np.all(ta >= 0.0)


def test_absorption_deletes_the_photon_and_kicks_the_electron(oracle):
    s = ac.build(oracle)
    el0, ph0, _, _ = ac.snapshot(s)
    n = ph0["cell"].size
    rng = np.random.default_rng(1)
    ta0 = np.where(rng.random(n) < 0.5, rng.uniform(0.0, 0.3, n), 50.0)
    ac.set_depths(s, ta0, np.full(n, np.inf))
    s.absorb(0.0)
    el1, ph1, ta, ts = ac.snapshot(s)
    gone = n - ph1["cell"].size
    assert 20 < gone < n
    assert np.all(ts == np.inf)
    assert np.all(ta >= 0.0)                                                  # whoever ran out was absorbed
    p0, p1 = _momentum(el0, ph0), _momentum(el1, ph1)
    assert np.all(np.abs(p1 - p0) <= 1e-12 * np.abs(p0).max())
    kicked = el0["ux"] != el1["ux"]
    assert 0 < kicked.sum() <= gone
    assert np.allclose(el1["gamma"], np.sqrt(1.0 + el1["ux"] ** 2 + el1["uy"] ** 2 + el1["uz"] ** 2), rtol=1e-15)
    assert not kicked[el0["cell"] == el0["cell"].max()].any()                 # interactions.rs:216-222


def test_stimulated_emission_adds_a_copy_and_takes_it_from_the_electron(oracle):
    s = ac.build(oracle)
    el0, ph0, _, _ = ac.snapshot(s)
    n = ph0["cell"].size
    rng = np.random.default_rng(2)
    ts0 = rng.uniform(0.0, 0.3, n)
    t = 7.0e-17
    ac.set_depths(s, np.full(n, np.inf), ts0)
    s.absorb(t)
    el1, ph1, ta, ts = ac.snapshot(s)
    added = ph1["cell"].size - n
    assert added > 20
    p0, p1 = _momentum(el0, ph0), _momentum(el1, ph1)
    assert np.all(np.abs(p1 - p0) <= 1e-12 * np.abs(p0).max())
    new = ph1["birth_time"] == t                                              # at_time(t); the uploaded ones were born at -inf
    assert new.sum() == added
    assert np.all(np.isin(ph1["weight"][new], el0["weight"]))                 # with_weight(e.weight())
    assert np.all(np.isfinite(ta[new]) & (ta[new] > 0.0) & np.isfinite(ts[new]) & (ts[new] > 0.0))     # with_optical_depths
    assert np.all(ts >= 0.0)                                                  # reset after the event


def test_flags_and_filters(oracle):
    s = ac.build(oracle, flags=FLAG_NO_STIMULATED_EMISSION)
    n = s.count(PHOTON)
    ac.set_depths(s, np.full(n, 50.0), np.full(n, 1.0e-6))
    s.absorb(0.0)
    _, _, ta, ts = ac.snapshot(s)
    assert np.all(ts == 1.0e-6) and np.any(ta < 50.0)                         # cfg!(not(feature = "no_stimulated_emission"))
    # stop_time: uploaded photons were "born" at -inf, so any finite limit switches them off (interactions.rs:181-185)
    s = ac.build(oracle)
    ac.set_depths(s, np.full(n, 50.0), np.full(n, 50.0))
    s.absorb(0.0, stop_time=1.0)
    _, ph, ta, ts = ac.snapshot(s)
    assert np.all(ta == 50.0) and np.all(ts == 50.0)
    # max_displacement: only photons that stayed on the axis (none here: every one has some k_perp) take part (:188-192)
    r = np.hypot(ph["y"], ph["z"])
    cut = np.median(r)
    s.absorb(0.0, max_displacement=cut)
    _, _, ta, ts = ac.snapshot(s)
    assert np.all(ta[r > cut] == 50.0) and np.any(ta[r <= cut] < 50.0)


def test_empty_populations(oracle):
    from opal_b200.abi import Slab
    s = Slab(oracle, ac.NX, 0.0, ac.DX, ac.DT)
    s.absorb(0.0)                                                             # interactions.rs:148-150
    assert s.count(PHOTON) == 0 and s.count(ELECTRON) == 0

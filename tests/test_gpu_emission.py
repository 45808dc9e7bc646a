"""QED-on parity: the random streams of the reference depend on thread count and sort
order, so photon emission is compared STATISTICALLY -- yield, energy spectrum, angle spectrum and
electron energy loss of the CUDA path against the CPU oracle at equal N, over several seeds."""
import numpy as np
import pytest
from scipy import stats

from opal_b200.abi import BDY_INTERNAL, ELECTRON, PHOTON, Slab
from opal_b200.constants import CRITICAL_FIELD, SPEED_OF_LIGHT
import pic_cases as pc

pytestmark = pytest.mark.gpu


def _run(abi, seed, chi0, nsteps=120, n=20000, filters=(None, None, None), flags=0):
    nx, dx = 256, 1.0e-8
    s = Slab(abi, nx, 0.0, dx, seed=seed, flags=flags)
    gamma0 = 1000.0
    e0 = chi0 * CRITICAL_FIELD / (2.0 * gamma0)          # head-on plane wave: chi ~ 2 gamma E / E_crit
    f = s.fields()
    f[:, 6] = e0                                          # Ey
    f[:, 10] = e0 / SPEED_OF_LIGHT                        # Bz  (wave travelling towards +x)
    s.set_fields(f)
    rng = np.random.default_rng(1000 + seed)
    cell = rng.integers(0, nx, n)
    ux = -gamma0 * (1.0 + 0.01 * rng.standard_normal(n))  # colliding_beams.yaml:19
    tau = rng.exponential(1.0, n)
    s.upload(ELECTRON, cell, rng.random(n), ux, np.zeros(n), np.zeros(n), np.ones(n), tau)
    t = 0.0
    for _ in range(nsteps):
        s.advance(ELECTRON)
        s.advance(PHOTON)
        s.emit_radiation(t, *filters)
        t += s.dt
    ph = s.download(PHOTON)
    el = s.download(ELECTRON)
    s.close()
    return ph, el


@pytest.mark.parametrize("chi0", [0.05, 0.8])
def test_photon_yield_and_spectra(gpu, oracle, chi0):
    yields_g, yields_o, eg, eo, ag, ao, gg, go = [], [], [], [], [], [], [], []
    for seed in range(5):
        pg, elg = _run(gpu, seed, chi0)
        po, elo = _run(oracle, seed + 50, chi0)
        yields_g.append(pg["cell"].size); yields_o.append(po["cell"].size)
        eg.append(pg["gamma"]); eo.append(po["gamma"])                    # photon energy / mc^2
        kperp_g = np.hypot(pg["uy"], pg["uz"]); kperp_o = np.hypot(po["uy"], po["uz"])
        ag.append(np.arctan2(kperp_g, -pg["ux"])); ao.append(np.arctan2(kperp_o, -po["ux"]))
        gg.append(elg["gamma"]); go.append(elo["gamma"])
    ng, no = float(np.sum(yields_g)), float(np.sum(yields_o))
    assert ng > 2.0e4 and no > 2.0e4
    # yields: two independent Poisson-like counts
    assert abs(ng - no) < 5.0 * np.sqrt(ng + no), (yields_g, yields_o)
    eg, eo, ag, ao = map(np.concatenate, (eg, eo, ag, ao))
    # two-sample KS on energy and polar-angle spectra
    assert stats.ks_2samp(np.log(eg), np.log(eo)).pvalue > 1.0e-3
    assert stats.ks_2samp(ag, ao).pvalue > 1.0e-3
    # radiated energy == electron energy loss statistically: mean final gamma agrees within 5 sigma
    gg, go = np.concatenate(gg), np.concatenate(go)
    se = np.sqrt(gg.var() / gg.size + go.var() / go.size)
    assert abs(gg.mean() - go.mean()) < 5.0 * se


def test_emission_filters_and_bookkeeping(gpu, oracle):
    """photons inherit (cell, x, weight) and birth time; the angle/energy filters of
    interactions.rs:74-95 only decide what is stored (the recoil always applies)"""
    # the crossed field turns the beam by ~12 mrad per step: a 0.1 rad cone keeps roughly the first 8 steps
    filt = (1.0, 0.1, None)         # >= 1 MeV, within 0.1 rad of -x
    pg, elg = _run(gpu, 3, 0.5, nsteps=30, filters=filt)
    po, elo = _run(oracle, 4, 0.5, nsteps=30, filters=filt)
    pa, ela = _run(gpu, 3, 0.5, nsteps=30)
    for p in (pg, po):
        assert np.all(p["gamma"] * 0.510999 >= 1.0)
        ang = np.arccos(-p["ux"] / p["gamma"])
        assert np.all(ang <= 0.1)
        assert np.all((p["birth_time"] >= 0.0) & np.isfinite(p["birth_time"]))
        assert np.all(p["weight"] == 1.0)
    assert 0 < pg["cell"].size < pa["cell"].size
    assert abs(pg["cell"].size - po["cell"].size) < 6.0 * np.sqrt(pg["cell"].size + po["cell"].size)
    # with / without filters the electrons recoil the same way (statistically: the slot a mover lands in,
    # and with it its Philox counter, depends on atomic ordering -- like the reference's thread-dependent streams)
    se = np.sqrt(elg["gamma"].var() / elg["gamma"].size + ela["gamma"].var() / ela["gamma"].size)
    assert abs(elg["gamma"].mean() - ela["gamma"].mean()) < 5.0 * se


def _run_fused(abi, monkeypatch, burst, chi0=0.5, n=40000, chunks=(37, 6, 50)):
    """the same beam through opal_run_steps with emission on (main.rs:238-267), several calls of uneven length"""
    monkeypatch.setenv("OPAL_BURST", burst)
    nx, dx = 256, 1.0e-8
    s = Slab(abi, nx, 0.0, dx, seed=7)
    gamma0 = 1000.0
    e0 = chi0 * CRITICAL_FIELD / (2.0 * gamma0)
    f = s.fields()
    f[:, 6] = e0
    f[:, 10] = e0 / SPEED_OF_LIGHT
    s.set_fields(f)
    rng = np.random.default_rng(4242)
    w = 1.0 + 1.0e-6 * np.arange(n)                                   # unique weights: photons inherit them (electron.rs:243)
    s.upload(ELECTRON, rng.integers(0, nx, n), rng.random(n), -gamma0 * (1.0 + 0.01 * rng.standard_normal(n)), np.zeros(n), np.zeros(n), w,
             rng.exponential(1.0, n))
    t = 0.0
    for k in chunks:
        s.run_steps(k, t, current_deposition=False, photon_emission=True, emit_filters=(None, None, None))
        t += k * s.dt
    ph, el, bursts = s.download(PHOTON), s.download(ELECTRON), s.burst_steps()
    s.close()
    return ph, el, bursts


def test_emission_inside_graph_replayed_bursts_matches_the_step_by_step_path(gpu, monkeypatch):
    """opal_run_steps with emission on runs as replayed CUDA graphs too (photon push, the push-built emitter list, the
    device-side room check of the photon store, k_emit reading time and call index from the device-side step counter).
    The draws are keyed by (seed, call index, bucket slot) and the slots of cell-crossing particles are handed out by
    atomics, so two runs are two realisations: yields and spectra are compared statistically; the times of birth -- the
    device-side step counter at work -- must be the very same set of values main.rs:266 accumulates."""
    pa, ea, ba = _run_fused(gpu, monkeypatch, "1")
    pb, eb, bb = _run_fused(gpu, monkeypatch, "0")
    assert ba > 60 and bb == 0
    na, nb = pa["cell"].size, pb["cell"].size
    assert na > 10000 and abs(na - nb) < 4.0 * np.sqrt(na + nb), (na, nb)
    ua, ub = np.unique(pa["birth_time"]), np.unique(pb["birth_time"])
    assert ua.size == ub.size == 93 and np.array_equal(ua, ub)           # every one of the 37 + 6 + 50 steps emitted, at the same t
    assert stats.ks_2samp(np.log(pa["gamma"]), np.log(pb["gamma"])).pvalue > 1.0e-3
    assert stats.ks_2samp(pa["x"] + pa["cell"], pb["x"] + pb["cell"]).pvalue > 1.0e-3
    # every photon carries the weight of the electron that emitted it (electron.rs:243); no electron is lost
    assert np.isin(pa["weight"], ea["weight"]).all() and ea["weight"].size == eb["weight"].size == 40000
    ga, gb = ea["gamma"], eb["gamma"]
    assert abs(ga.mean() - gb.mean()) < 4.0 * np.sqrt(ga.var() / ga.size + gb.var() / gb.size)

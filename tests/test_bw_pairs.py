"""EXTENSION: nonlinear Breit-Wheeler pair creation (OPAL_FLAG_PAIR_CREATION, include/opal_b200.h).  north_star names it under
src/qed, but the reference implements none of it (only the unused optical depth Photon.tau, photon.rs:27,126-132): PARITY
UNPINNED -- there is nothing of opal's to compare with.  What is checked instead:
  * the rate function F(chi) = chi T(chi) and the energy-share spectrum dW/df against direct quadrature of the published
    formula (Baier-Katkov / Ritus; tools/gen_bw_tables.py) and against the published asymptotes
    F -> 3 sqrt(3)/(16 sqrt(2)) chi exp(-8/(3 chi)) for chi << 1 and F -> 0.38 chi^(2/3) for chi >> 1;
  * yield: photons in a constant crossed field convert as N0 (1 - exp(-W t)), W = alpha_f F / (tau_C k0);
  * bookkeeping per event: u_electron + u_positron = k_photon exactly, both inherit the photon's weight and position;
  * the CUDA kernel (in k_photon_push) against the CPU restatement statistically; flag off => nothing changes."""
import math
import sys
import pathlib

import numpy as np
import pytest
from scipy import stats

from opal_b200.abi import ELECTRON, FLAG_PAIR_CREATION, ION, PHOTON, Slab
from opal_b200.constants import CRITICAL_FIELD, SPEED_OF_LIGHT

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
ALPHA_FINE, COMPTON_TIME = 7.29735257e-3, 1.28808867e-21      # constants.rs
LN_CHI0, DLN = math.log(0.05), (math.log(500.0) - math.log(0.05)) / 80


def _f_and_share(abi, tag):
    import gen_bw_tables as gen
    # the rate function against quadrature of the published spectrum (between table nodes: linear interpolation of ln F)
    for chi, tol in ((0.3, 2e-2), (1.0, 5e-3), (3.7, 2e-3), (40.0, 1e-3), (333.0, 1e-3)):
        assert abi.bw_f(chi)[0] == pytest.approx(gen.big_f(chi), rel=tol), (tag, chi)
    # published asymptotes, at table nodes near the two ends and beyond them (where the functions switch to the asymptotes)
    small = lambda c: 3.0 * math.sqrt(3.0) / (16.0 * math.sqrt(2.0)) * c * math.exp(-8.0 / (3.0 * c))
    for chi in (math.exp(LN_CHI0), math.exp(LN_CHI0 + 2 * DLN), 0.03, 0.01):
        assert abi.bw_f(chi)[0] == pytest.approx(small(chi), rel=2e-2), (tag, chi)
    for chi in (math.exp(LN_CHI0 + 79 * DLN), 500.0, 5000.0):
        assert abi.bw_f(chi)[0] == pytest.approx(0.3796 * chi ** (2.0 / 3.0), rel=4e-2), (tag, chi)
    assert abi.bw_f(0.0)[0] == 0.0
    # the share spectrum: samples against the numerically integrated CDF of dW/df
    rng = np.random.default_rng(3)
    for chi in (0.7, 8.0):
        n = 40000
        f = abi.bw_share(np.full(n, chi), rng.random(n), rng.random(n))
        assert f.min() > 0.0 and f.max() < 1.0 and abs(f.mean() - 0.5) < 4.0 * f.std() / math.sqrt(n)
        grid = np.linspace(0.0, 1.0, 401)
        dens = np.array([gen.spectrum(chi, x) for x in grid])
        cdf = np.concatenate([[0.0], np.cumsum(0.5 * (dens[1:] + dens[:-1]) * np.diff(grid))])
        cdf /= cdf[-1]
        assert stats.kstest(f, lambda x: np.interp(x, grid, cdf)).pvalue > 1e-3, (tag, chi)
    # r2 decides which lepton takes the smaller share
    assert (abi.bw_share(np.full(100, 2.0), rng.random(100), np.full(100, 0.25)) <= 0.5).all()
    assert (abi.bw_share(np.full(100, 2.0), rng.random(100), np.full(100, 0.75)) >= 0.5).all()


def _beam_in_crossed_field(abi, flags, chi=2.0, k0=2000.0, n=30000, steps=60, seed=11):
    nx, dx = 128, 1.0e-8
    s = Slab(abi, nx, 0.0, dx, seed=seed, flags=flags)
    e0 = chi * CRITICAL_FIELD / (2.0 * k0)                 # head-on plane wave: chi = 2 k0 E / E_crit
    f = s.fields()
    f[:, 6] = e0
    f[:, 10] = e0 / SPEED_OF_LIGHT
    s.set_fields(f)
    rng = np.random.default_rng(100 + seed)
    w = 1.0 + 1.0e-6 * np.arange(n)                        # unique: identifies the photon a lepton came from
    s.upload(PHOTON, rng.integers(0, nx, n), rng.random(n), np.full(n, -k0), np.zeros(n), np.zeros(n), w, rng.exponential(1.0, n))
    for _ in range(steps):
        s.advance(PHOTON)
    out = dict(ph=s.download(PHOTON), el=s.download(ELECTRON), po=s.download(ION), pairs=s.pairs_created(), dt=s.dt, k0=k0, n=n, steps=steps)
    s.close()
    return out


def _check_bookkeeping(r, abi):
    el, po, ph = r["el"], r["po"], r["ph"]
    assert el["weight"].size == po["weight"].size == r["pairs"] and r["pairs"] + ph["weight"].size == r["n"]      # a photon or its pair, never both
    ie, ip = np.argsort(el["weight"]), np.argsort(po["weight"])
    assert np.array_equal(el["weight"][ie], po["weight"][ip])
    assert not np.isin(el["weight"], ph["weight"]).any()
    assert np.array_equal(el["cell"][ie], po["cell"][ip]) and np.array_equal(el["x"][ie], po["x"][ip])
    # momentum: exact
    assert np.abs(el["ux"][ie] + po["ux"][ip] + r["k0"]).max() <= 1e-12 * r["k0"]
    assert np.abs(el["uy"]).max() == 0.0 and np.abs(po["uz"]).max() == 0.0
    # energy: gamma_e + gamma_p = k0 (1 + O(1/k0^2)); the positron is an "ion" of mass m_e: gamma = 1 + gamma_m1
    assert np.abs(el["gamma"][ie] + po["gamma"][ip] - r["k0"]).max() < 2.0 / r["k0"] * 10
    # yield: N0 (1 - exp(-W t)); every photon carries its own Exp(1) depth, so conversions are independent Bernoulli trials
    w_rate = ALPHA_FINE * abi.bw_f(2.0)[0] / (COMPTON_TIME * r["k0"])
    p = 1.0 - math.exp(-w_rate * r["dt"] * r["steps"])
    assert 0.3 < p < 0.95
    assert abs(r["pairs"] - r["n"] * p) < 4.0 * math.sqrt(r["n"] * p * (1.0 - p)), (r["pairs"], r["n"] * p)
    return el["ux"][ie] / -r["k0"]                       # the electron's share f


def test_rate_and_share_of_the_cpu_restatement(oracle):
    _f_and_share(oracle, "oracle")


def test_pair_creation_in_a_crossed_field_cpu_restatement(oracle):
    r = _beam_in_crossed_field(oracle, FLAG_PAIR_CREATION)
    f = _check_bookkeeping(r, oracle)
    assert abs(f.mean() - 0.5) < 4.0 * f.std() / math.sqrt(f.size)
    off = _beam_in_crossed_field(oracle, 0, steps=5)
    assert off["pairs"] == 0 and off["ph"]["weight"].size == off["n"] and off["el"]["weight"].size == 0


@pytest.mark.gpu
def test_rate_and_share_on_the_gpu(gpu, oracle):
    _f_and_share(gpu, "gpu")
    chi = np.exp(np.linspace(math.log(0.02), math.log(2000.0), 500))
    assert np.abs(gpu.bw_f(chi) / oracle.bw_f(chi) - 1.0).max() < 1e-12
    rng = np.random.default_rng(8)
    r1, r2 = rng.random(500), rng.random(500)
    assert np.abs(gpu.bw_share(chi, r1, r2) - oracle.bw_share(chi, r1, r2)).max() < 1e-12


@pytest.mark.gpu
def test_pair_creation_in_a_crossed_field_gpu_vs_cpu_restatement(gpu, oracle):
    g = _beam_in_crossed_field(gpu, FLAG_PAIR_CREATION)
    o = _beam_in_crossed_field(oracle, FLAG_PAIR_CREATION, seed=12)
    fg, fo = _check_bookkeeping(g, gpu), _check_bookkeeping(o, oracle)
    assert abs(g["pairs"] - o["pairs"]) < 4.0 * math.sqrt(g["pairs"] + o["pairs"])
    assert stats.ks_2samp(fg, fo).pvalue > 1e-3
    # the new electrons carry fresh Exp(1) optical depths; the photons that are left carry depths that ran down but not out
    assert stats.kstest(g["el"]["tau"], "expon").pvalue > 1e-3
    assert (g["ph"]["tau"] >= 0.0).all()


@pytest.mark.gpu
def test_flag_off_changes_nothing(gpu):
    """without OPAL_FLAG_PAIR_CREATION the photon advance neither touches the optical depth nor creates anything"""
    r = _beam_in_crossed_field(gpu, 0, steps=7, n=5000)
    assert r["pairs"] == 0 and r["el"]["weight"].size == 0 and r["po"]["weight"].size == 0
    rng = np.random.default_rng(100 + 11)
    rng.integers(0, 128, 5000); rng.random(5000)
    assert np.array_equal(np.sort(r["ph"]["tau"]), np.sort(rng.exponential(1.0, 5000)))


@pytest.mark.gpu
def test_set_ion_is_refused_with_pair_creation(gpu):
    from opal_b200.abi import OpalError
    s = Slab(gpu, 16, 0.0, 1.0e-8, flags=FLAG_PAIR_CREATION)
    with pytest.raises(OpalError):
        s.set_ion(6.0, 12.0)
    s.close()

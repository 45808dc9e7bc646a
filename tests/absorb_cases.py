"""Photon absorption / stimulated emission scenarios (particle/interactions.rs:145) shared by the oracle's own tests
and the CUDA parity tests.

A short time step (0.1 dx / c) and particles placed mid-cell keep every particle in its cell over the one advance that
gives electrons and photons their chi, so the electrons of a cell sit in the same order in the reference's sorted store
and in the CUDA library's cell buckets: the pair loop's "first electron whose partial probability exhausts the optical
depth" is then the same electron on both sides."""
import numpy as np

from opal_b200.abi import ELECTRON, PHOTON, Slab
from opal_b200.constants import CRITICAL_FIELD, SPEED_OF_LIGHT

NX, DX = 24, 1.0e-8
DT = 0.1 * DX / SPEED_OF_LIGHT
GAMMA0 = 1000.0
# macro-electron weight that makes one photon-electron pair worth an optical depth of order 1e-2..1
E_WEIGHT = 2.0e27


def build(abi, flags=0, seed=5, n_e=400, n_ph=300, chi0=1.0, last_cell=20):
    """fields of a plane wave meeting the beam head on; electrons in cells 2..last_cell, photons in cells 0..NX-1;
    one advance of both species (no cell changes) so that chi is set"""
    s = Slab(abi, NX, 0.0, DX, DT, flags=flags, seed=seed)
    e0 = chi0 * CRITICAL_FIELD / (2.0 * GAMMA0)
    f = s.fields()
    f[:, 6] = e0
    f[:, 10] = e0 / SPEED_OF_LIGHT
    s.set_fields(f)
    rng = np.random.default_rng(seed)
    cell = np.sort(rng.integers(2, last_cell + 1, n_e))
    ux = -GAMMA0 * (1.0 + 0.2 * rng.random(n_e))
    uy, uz = 3.0 * rng.standard_normal(n_e), 3.0 * rng.standard_normal(n_e)
    w = E_WEIGHT * (0.5 + rng.random(n_e))
    s.upload(ELECTRON, cell, 0.3 + 0.4 * rng.random(n_e), ux, uy, uz, w)
    pcell = np.sort(rng.integers(0, NX, n_ph))
    k0 = GAMMA0 * 10.0 ** rng.uniform(-3.0, 0.2, n_ph)              # some above the electrons' energy: no stimulated emission there
    th = 2.0e-3 * rng.standard_normal(n_ph)
    ph = rng.uniform(0.0, 2.0 * np.pi, n_ph)
    s.upload(PHOTON, pcell, 0.3 + 0.4 * rng.random(n_ph), -k0 * np.cos(th), k0 * np.sin(th) * np.cos(ph), k0 * np.sin(th) * np.sin(ph),
             0.01 * E_WEIGHT * (1.0 + rng.random(n_ph)))       # an absorbed photon moves its electron by ~1% of k
    s.advance(ELECTRON)
    s.advance(PHOTON)
    return s


def photon_key(ph):
    """canonical photon order: by cell, position, momentum (then birth time and weight, which tell a stimulated copy from
    its original); the two libraries keep the photons of a cell in different orders, and the pair loop does not care"""
    return np.lexsort((ph["weight"], ph["birth_time"], ph["uz"], ph["uy"], ph["ux"], ph["x"], ph["cell"]))


def snapshot(s):
    """electrons in store order; photons and their (tau_abs, tau_st) in canonical order"""
    el, ph = s.download(ELECTRON), s.download(PHOTON)
    ta, ts = s.photon_pair_depths()
    o = photon_key(ph)
    return el, {k: v[o] for k, v in ph.items()}, ta[o], ts[o]


def set_depths(s, tau_abs, tau_st):
    """tau_abs / tau_st given in canonical photon order"""
    o = photon_key(s.download(PHOTON))
    a, b = np.empty(o.size), np.empty(o.size)
    a[o], b[o] = tau_abs, tau_st
    s.set_photon_pair_depths(a, b)

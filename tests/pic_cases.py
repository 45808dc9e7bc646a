"""Seeded PIC test cases and comparison helpers shared by the oracle/GPU parity tests.

Initial particles are generated ONCE with numpy and the same arrays are uploaded to both the
oracle and the CUDA library (the reference's loader RNG is unpinned, so deterministic
parity starts from identical arrays).  Every particle carries a unique weight so particles can be
matched one-to-one after the (order-unspecified) per-cell sort.
"""
import math

import numpy as np

from opal_b200.abi import (BDY_ABSORBING, BDY_INTERNAL, BDY_LASER, ELECTRON, ION, PHOTON, Slab)
from opal_b200.constants import ELECTRON_MASS, SPEED_OF_LIGHT

C = SPEED_OF_LIGHT
FLOAT_TOL = 1.0e-10     # north_star: fields, currents, momenta within 1e-10 relative per step


def two_stream_particles(nx, npc, dx, seed=1, ne=20.0, thermal=0.0):
    """examples/two_stream.yaml:14-20 shaped population: counter-streaming cold beams."""
    rng = np.random.default_rng(seed)
    n = nx * npc
    cell = np.repeat(np.arange(nx, dtype=np.int64), npc)
    x = rng.random(n)
    drift_u = 2.5e-24 / (ELECTRON_MASS * C)
    urand, nrand = rng.random(n), rng.standard_normal(n)
    ux = drift_u * (1.0 + 0.001 * nrand) * (2.0 * np.floor(urand - 0.5) + 1.0)
    uy = thermal * rng.standard_normal(n)
    uz = thermal * rng.standard_normal(n)
    w0 = ne * dx / npc
    w = w0 * (1.0 + 1.0e-9 * np.arange(n))       # unique ids
    return dict(cell=cell, x=x, ux=ux, uy=uy, uz=uz, weight=w)


def thermal_particles(nx, npc, dx, u_th, seed=2, ne=20.0, cells=None):
    rng = np.random.default_rng(seed)
    cells = np.arange(nx, dtype=np.int64) if cells is None else np.asarray(cells, dtype=np.int64)
    n = cells.size * npc
    cell = np.repeat(cells, npc)
    w0 = ne * dx / npc
    return dict(cell=cell, x=rng.random(n), ux=u_th * rng.standard_normal(n), uy=u_th * rng.standard_normal(n),
                uz=u_th * rng.standard_normal(n), weight=w0 * (1.0 + 1.0e-9 * np.arange(n)))


def upload(slab, species, p, tau=None):
    slab.upload(species, p["cell"], p["x"], p["ux"], p["uy"], p["uz"], p["weight"], tau)


def canon(d):
    """sort a downloaded population by (cell, weight id)"""
    order = np.lexsort((d["weight"], d["cell"]))
    return {k: v[order] for k, v in d.items()}


def rel_inf(a, b):
    """norm-wise relative error  ||a-b||_inf / ||b||_inf  (SURVEY 8c parity criterion)"""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    if a.shape != b.shape:
        return math.inf
    fin = np.isfinite(b)
    if not np.array_equal(fin, np.isfinite(a)):
        return math.inf
    if not np.array_equal(a[~fin], b[~fin], equal_nan=True):
        return math.inf
    if not fin.any():
        return 0.0
    scale = np.max(np.abs(b[fin]))
    diff = np.max(np.abs(a[fin] - b[fin]))
    return 0.0 if diff == 0.0 else diff / (scale if scale > 0 else 1.0)


FIELD_NAMES = ["x", "rho", "jx", "jy", "jz", "Ex", "Ey", "Ez", "Bx", "By", "Bz"]


def compare_fields(fa, fb, tol=FLOAT_TOL, skip=()):
    worst = {}
    for k, name in enumerate(FIELD_NAMES):
        if name in skip:
            continue
        worst[name] = rel_inf(fa[:, k], fb[:, k])
    bad = {k: v for k, v in worst.items() if not v <= tol}
    assert not bad, f"field mismatch beyond {tol}: {bad}"
    return worst


def compare_particles(da, db, fields=("x", "ux", "uy", "uz"), tol=FLOAT_TOL):
    a, b = canon(da), canon(db)
    # integer / index work: bit-exact
    assert a["cell"].shape == b["cell"].shape, (a["cell"].shape, b["cell"].shape)
    assert np.array_equal(a["cell"], b["cell"]), "cell keys differ"
    assert np.array_equal(a["weight"], b["weight"]), "particle identity (weight) differs"
    worst = {f: rel_inf(a[f], b[f]) for f in fields}
    bad = {k: v for k, v in worst.items() if not v <= tol}
    assert not bad, f"particle mismatch beyond {tol}: {bad}"
    return worst


def step_by_calls(slab, t, deposition=True, laser=None, species=(ELECTRON, ION, PHOTON), emission=None):
    """one iteration of main.rs:238-267 through the call-by-call API"""
    for sp in species:
        slab.advance(sp)
    if emission is not None:
        slab.emit_radiation(t, *emission)
    if deposition:
        slab.clear()
        for sp in species:
            if sp != PHOTON:
                slab.deposit(sp)
    if laser is None:
        slab.synchronize(t, 0.0, 0.0)
    else:
        x2 = slab.laser_x()
        slab.synchronize(t, laser[0](t, x2), laser[1](t, x2))
    slab.field_advance()
    return t + slab.dt

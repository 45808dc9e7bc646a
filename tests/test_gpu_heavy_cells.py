"""The lean strip kernel hands a cell holding more than OPAL_HEAVY_MIN particles to a whole block (k_push_lean_fast<.., HEAVY>,
opal_push_lean.cuh): its warps share the cell's residents and arrivals and hand out strip and free slots through counters
in shared memory.  Same particles, same fields, same leaver counts as the oracle -- whichever launch a cell goes through,
also when it changes sides from one step to the next, and with the launch switching itself on in the middle of a burst."""
import numpy as np
import pytest

from opal_b200.abi import BDY_ABSORBING, BDY_INTERNAL, BDY_LASER, ELECTRON, FLAG_LEAN_STATE, ION, Slab
import pic_cases as pc

pytestmark = pytest.mark.gpu


def _energies_agree(a, b, tol=1e-10):
    ea, eb = a.energy(), b.energy()
    for k in range(4):
        assert abs(ea[k] - eb[k]) <= tol * abs(eb[k]) + 1e-300, (k, ea, eb)


def _pile(nx, dx, seed, u_th, heavy_cells, n_heavy, n_light, ne=20.0):
    """a dense pile (n_heavy per cell) inside a thin plasma (n_light per cell): equal weights, so the pile is a real density peak"""
    light = np.setdiff1d(np.arange(nx), heavy_cells)
    p1 = pc.thermal_particles(nx, n_heavy, dx, u_th, seed=seed, cells=heavy_cells, ne=ne * n_heavy / n_light)
    p2 = pc.thermal_particles(nx, n_light, dx, u_th, seed=seed + 1, cells=light, ne=ne)
    p2["weight"] = p2["weight"] * 1.000123
    return {k: np.concatenate([p1[k], p2[k]]) for k in p1}


@pytest.mark.parametrize("force", [0, 1])
def test_dense_pile_matches_oracle(gpu, oracle, monkeypatch, force):
    """periodic slab, a pile of 3000-particle cells drifting and spreading through 150-particle cells: cells cross the
    threshold in both directions; force=0: the block-per-cell launch switches itself on (the first chunk of steps still runs
    without it), force=1: on from the first step, with a threshold low enough that the thin cells take it too"""
    monkeypatch.setenv("OPAL_HEAVY_MIN", "100" if force else "512")
    monkeypatch.setenv("OPAL_HEAVY_FACTOR", "0")              # (the pile is only six times the slab's average)
    monkeypatch.setenv("OPAL_HEAVY_FORCE", str(force))
    nx, dx = 96, 500.0
    a = Slab(gpu, nx, 0.0, dx, flags=FLAG_LEAN_STATE)
    b = Slab(oracle, nx, 0.0, dx)
    p = _pile(nx, dx, 11, 0.3, np.arange(40, 52), 3000, 150)
    p["ux"] = p["ux"] + 0.2                                   # the pile moves: new cells become heavy, old ones thin out
    f = b.fields()
    k = np.arange(f.shape[0])
    f[:, pc.FIELD_NAMES.index("By")] = 1.0e-7 * np.cos(0.05 * k)
    f[:, pc.FIELD_NAMES.index("Ey")] = 30.0 * np.sin(0.09 * k)
    for s_ in (a, b):
        s_.set_fields(f)
        pc.upload(s_, ELECTRON, p)
    monkeypatch.setenv("OPAL_HEAVY_MIN", "0")                 # a third slab that never uses the block-per-cell launch
    a0 = Slab(gpu, nx, 0.0, dx, flags=FLAG_LEAN_STATE)
    a0.set_fields(f)
    pc.upload(a0, ELECTRON, p)
    for chunk in (1, 3, 12, 40):
        a.run_steps(chunk)
        a0.run_steps(chunk)
        b.run_steps(chunk)
        pc.compare_particles(a.download(ELECTRON), b.download(ELECTRON), fields=("x", "ux", "uy", "uz"))
        pc.compare_fields(a.fields(), b.fields())
        assert a.last_leavers(ELECTRON) == b.last_leavers(ELECTRON)
        _energies_agree(a, b)
    assert a.count(ELECTRON) == p["cell"].size
    pc.compare_particles(a0.download(ELECTRON), b.download(ELECTRON), fields=("x", "ux", "uy", "uz"))
    assert a.launch_count() > a0.launch_count() + 20          # the second launch per push did run (force=0: came on by itself)


def test_dense_pile_two_species_open_boundaries(gpu, oracle, monkeypatch):
    """laser + absorbing boundary, electrons and C6+ ions, the pile at the slab's edge cells too (leavers of a heavy edge cell are
    deleted and counted), general B_x rotation"""
    monkeypatch.setenv("OPAL_HEAVY_MIN", "200")
    monkeypatch.setenv("OPAL_HEAVY_FACTOR", "0")
    monkeypatch.setenv("OPAL_HEAVY_FORCE", "1")
    nx, dx = 80, 1.0e-8
    a = Slab(gpu, nx, 0.0, dx, None, BDY_LASER, BDY_ABSORBING, flags=FLAG_LEAN_STATE)
    b = Slab(oracle, nx, 0.0, dx, None, BDY_LASER, BDY_ABSORBING)
    heavy = np.concatenate([np.arange(0, 3), np.arange(30, 36), np.arange(nx - 2, nx)])
    pe = _pile(nx, dx, 21, 0.4, heavy, 1500, 60, ne=1.0e20)
    pi = _pile(nx, dx, 31, 0.05, heavy, 900, 40, ne=1.0e20 / 6.0)
    f = b.fields()
    k = np.arange(f.shape[0])
    f[:, pc.FIELD_NAMES.index("Bx")] = 3.0e-7
    f[:, pc.FIELD_NAMES.index("Bz")] = -1.5e-7 * np.sin(0.05 * k)
    f[:, pc.FIELD_NAMES.index("Ex")] = 10.0 * np.sin(0.11 * k)
    for s_ in (a, b):
        s_.set_fields(f)
        s_.set_ion(6.0, 12.0)
        pc.upload(s_, ELECTRON, pe)
        pc.upload(s_, ION, pi)
    for chunk in (1, 2, 15):
        a.run_steps(chunk)
        b.run_steps(chunk)
        for sp in (ELECTRON, ION):
            assert a.last_leavers(sp) == b.last_leavers(sp)
            assert a.count(sp) == b.count(sp)
            pc.compare_particles(a.download(sp), b.download(sp), fields=("x", "ux", "uy", "uz"))
        pc.compare_fields(a.fields(), b.fields())
        _energies_agree(a, b)
    assert a.count(ELECTRON) < pe["cell"].size                # some electrons did leave through the open edges


def test_one_huge_cell(gpu, oracle, monkeypatch):
    """20 000 particles in a single cell (more free slots per step than the block's free list holds: the rest stay free, the cell
    compacts itself) next to empty cells that only gain"""
    monkeypatch.setenv("OPAL_HEAVY_MIN", "512")
    monkeypatch.setenv("OPAL_HEAVY_FORCE", "0")
    nx, dx = 48, 500.0
    a = Slab(gpu, nx, 0.0, dx, flags=FLAG_LEAN_STATE)
    b = Slab(oracle, nx, 0.0, dx)
    p = pc.thermal_particles(nx, 20000, dx, 0.5, seed=3, cells=np.array([20]))
    pc.upload(a, ELECTRON, p)
    pc.upload(b, ELECTRON, p)
    for chunk in (2, 6, 20):
        a.run_steps(chunk)
        b.run_steps(chunk)
        pc.compare_particles(a.download(ELECTRON), b.download(ELECTRON), fields=("x", "ux", "uy", "uz"))
        pc.compare_fields(a.fields(), b.fields())
        assert a.last_leavers(ELECTRON) == b.last_leavers(ELECTRON)
    assert a.count(ELECTRON) == 20000


def test_wide_strips_bucket_overflow_while_a_cell_compacts(gpu, oracle, monkeypatch):
    """strips wide enough to carry hundreds of movers per cell and step (OPAL_CELL_MOVER_FRACTION): the thin cells ahead of the
    moving pile take in more arrivals than their buckets hold (the rest is parked for the host's re-layout) in the same step in
    which they close their free slots -- the slots the demand count then covers behind the live particles must read as free
    (they once held copies of the particles shifted down: 12 duplicates after 16 steps before the fix)"""
    monkeypatch.setenv("OPAL_HEAVY_MIN", "0")
    monkeypatch.setenv("OPAL_STRIP_WIDEN", "0")
    monkeypatch.setenv("OPAL_CELL_MOVER_FRACTION", "0.5")
    nx, dx = 96, 500.0
    a = Slab(gpu, nx, 0.0, dx, flags=FLAG_LEAN_STATE)
    b = Slab(oracle, nx, 0.0, dx)
    p = _pile(nx, dx, 11, 0.3, np.arange(40, 52), 3000, 150)
    p["ux"] = p["ux"] + 0.2
    pc.upload(a, ELECTRON, p)
    pc.upload(b, ELECTRON, p)
    for chunk in (1, 3, 12, 24):
        a.run_steps(chunk)
        b.run_steps(chunk)
        assert a.count(ELECTRON) == b.count(ELECTRON)
        pc.compare_particles(a.download(ELECTRON), b.download(ELECTRON), fields=("x", "ux", "uy", "uz"))
        pc.compare_fields(a.fields(), b.fields())
        assert a.last_leavers(ELECTRON) == b.last_leavers(ELECTRON)


def test_strips_widen_themselves(gpu, oracle, monkeypatch):
    """the same pile with the default (narrow) strips: most movers find their strip full, the host widens the strips between
    bursts (OPAL_STRIP_WIDEN, on by default) -- the movers waiting in the old strips are settled first -- and nothing changes"""
    monkeypatch.setenv("OPAL_HEAVY_MIN", "512")
    nx, dx = 96, 500.0
    a = Slab(gpu, nx, 0.0, dx, flags=FLAG_LEAN_STATE)
    b = Slab(oracle, nx, 0.0, dx)
    p = _pile(nx, dx, 13, 0.45, np.arange(20, 30), 2500, 120)
    pc.upload(a, ELECTRON, p)
    pc.upload(b, ELECTRON, p)
    for chunk in (2, 20, 20, 30):
        a.run_steps(chunk)
        b.run_steps(chunk)
        pc.compare_particles(a.download(ELECTRON), b.download(ELECTRON), fields=("x", "ux", "uy", "uz"))
        pc.compare_fields(a.fields(), b.fields())
        assert a.last_leavers(ELECTRON) == b.last_leavers(ELECTRON)
        _energies_agree(a, b)

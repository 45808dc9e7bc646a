"""N > 1 host logic on CPU: two gloo ranks, one oracle slab each, host-staged halo + migration
exchange in the reference's even/odd order (opal_b200/slabs.py), against ONE oracle slab covering
the whole domain.  Cell keys and population sizes are bit-exact; x, u and fields agree to 1e-10
(the two decompositions add ghost-cell currents in a different order)."""
import ctypes
import socket
import subprocess
import sys

import numpy as np
import pytest

import mslab_worker as w
import pic_cases as pc
from opal_b200.abi import ELECTRON, Slab
from opal_b200.slabs import GridDesign


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _run_two_ranks(case, tmp_path, balanced):
    port = _free_port()
    procs = [subprocess.Popen([sys.executable, w.__file__, str(r), "2", str(port), case, str(tmp_path), "1" if balanced else "0"])
             for r in range(2)]
    for p in procs:
        assert p.wait(timeout=300) == 0
    return [np.load(tmp_path / f"rank{r}.npz") for r in range(2)]


@pytest.mark.parametrize("case,balanced", [("periodic", False), ("laser", False), ("laser", True)])
def test_two_slabs_match_one(case, balanced, oracle, tmp_path):
    ranks = _run_two_ranks(case, tmp_path, balanced)
    g = w.global_case(case)
    one = Slab(oracle, g["nx"], g["xmin"], g["dx"], None, g["left"], g["right"])
    pc.upload(one, ELECTRON, g["particles"])
    laser = w.laser_fns(g["laser"]) if g["laser"] else None
    t = 0.0
    for _ in range(g["steps"]):
        t = pc.step_by_calls(one, t, laser=laser, species=(ELECTRON,))
    ref = one.download(ELECTRON)
    got = {k: np.concatenate([r["p_" + k] for r in ranks]) for k in ref}
    pc.compare_particles(got, ref, fields=("x", "prev_x", "ux", "uy", "uz", "gamma"))
    # interior cells of both slabs, stitched, against the single grid's interior
    nx = ranks[0]["nx"]
    assert int(nx.sum()) == g["nx"]
    if balanced:
        assert nx[0] != nx[1]
    stitched = np.concatenate([r["fields"][int(r["left_size"]): int(r["left_size"]) + int(n)] for r, n in zip(ranks, nx)])
    f1 = one.fields()
    pc.compare_fields(stitched, f1[one.left_size: one.left_size + g["nx"]])
    # energies reduced over ranks == single-slab energies
    e2, e1 = ranks[0]["energy"], np.array(one.energy())
    assert np.allclose(e2, e1, rtol=1e-10, atol=0.0)
    # something actually migrated between the slabs
    assert sum(int(r["leavers"].sum()) for r in ranks) > 0


def test_grid_design_matches_reference_rules():
    d = GridDesign.unbalanced(3, 1000, -5.0, 0.5)
    assert d.nx == [333, 333, 333] and d.offset == [0.0, 166.5, 333.0]        # nx mod np cells dropped (mod.rs:128)
    d = GridDesign.unbalanced(4, 20, 0.0, 1.0)
    assert d.nx == [8, 8, 8, 8]                                               # max(min_subsize) (mod.rs:128)
    d = GridDesign.balanced(2, 1000, 0.0, 1.0, lambda x: 1.0 if x < 250 else 0.0)
    assert sum(d.nx) == 1000 and d.nx[0] < 200                                # equal integral of ne, not equal cells
    d = GridDesign.balanced(4, 4000, 0.0, 1.0, lambda x: 1.0)
    assert sum(d.nx) == 4000 and all(abs(n - 1000) <= 10 for n in d.nx)


@pytest.mark.parametrize("case,world,balanced", [("periodic", 3, False), ("laser", 2, True), ("laser_sparse", 4, False)])
def test_inprocess_oracle_ring_matches_one_slab(case, world, balanced, oracle):
    """tests/ring_oracle.py (the checker of the multi-GPU runs): a ring of oracle slabs stepped rank after rank in one process
    equals one oracle slab covering the domain; the per-step leaver counts it reports are consistent with the particles that
    changed slab"""
    import ring_oracle
    g = w.global_case(case)
    if balanced:
        dens = ring_oracle.density_of(g["particles"], g["nx"])
        design = GridDesign.balanced(world, g["nx"], g["xmin"], g["dx"], lambda x: dens[min(int(round((x - g["xmin"]) / g["dx"])), g["nx"] - 1)],
                                     left=g["left"], right=g["right"])
        assert len(set(design.nx)) > 1
    else:
        design = GridDesign.unbalanced(world, g["nx"], g["xmin"], g["dx"], left=g["left"], right=g["right"])
    ring = ring_oracle.OracleRing(oracle, design)
    ring.upload(ELECTRON, g["particles"])
    one = Slab(oracle, g["nx"], g["xmin"], g["dx"], None, g["left"], g["right"])
    pc.upload(one, ELECTRON, g["particles"])
    laser = w.laser_fns(g["laser"]) if g["laser"] else None
    fields = ("cell", "x", "ux", "uy", "uz", "weight")
    t = 0.0
    total_gone = 0
    for k in range(g["steps"]):
        owner_before = {r: set(d["weight"].tolist()) for r, d in enumerate(ring.download(ELECTRON, ("cell", "weight")))} if k < 3 else None
        t1, gone = ring.step(t, laser=laser, species=(ELECTRON,))
        if owner_before is not None:     # first steps: count who changed slab by identity (unique weights)
            owner_after = {r: set(d["weight"].tolist()) for r, d in enumerate(ring.download(ELECTRON, ("cell", "weight")))}
            for r in range(world):
                left_r = len(owner_before[r] - owner_after[r])
                assert left_r == sum(gone[ELECTRON][r]), (r, left_r, gone[ELECTRON][r])
        total_gone += sum(sum(x) for x in gone[ELECTRON])
        pc.step_by_calls(one, t, laser=laser, species=(ELECTRON,))
        t = t1
    assert total_gone > 0
    got = {k: np.concatenate([d[k] for d in ring.download(ELECTRON, fields)]) for k in fields}
    pc.compare_particles(got, one.download(ELECTRON, fields=fields), fields=("x", "ux", "uy", "uz"))
    pc.compare_fields(np.concatenate(ring.interior_fields()), one.fields()[one.left_size: one.left_size + sum(design.nx)])
    ring.close()

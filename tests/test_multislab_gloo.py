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

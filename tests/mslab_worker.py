"""Worker for the world-size-2 gloo tests: one oracle slab per rank, host-staged ring exchange
(opal_b200.slabs.SlabRing).  Run as a spawned process; writes its final state to an .npz file."""
import ctypes
import os
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def global_case(case):
    import pic_cases as pc
    from opal_b200.abi import BDY_ABSORBING, BDY_INTERNAL, BDY_LASER
    if case == "periodic":
        nx, npc, dx = 96, 24, 500.0
        p = pc.two_stream_particles(nx, npc, dx, seed=5, thermal=0.25)
        return dict(nx=nx, dx=dx, xmin=0.0, left=BDY_INTERNAL, right=BDY_INTERNAL, particles=p, steps=40, laser=None)
    if case == "laser":
        nx, dx = 120, 2.0e-8
        p = pc.thermal_particles(nx, 12, dx, 0.4, seed=9, ne=5.0e26, cells=np.arange(30, 100))
        return dict(nx=nx, dx=dx, xmin=-1.0e-6, left=BDY_LASER, right=BDY_ABSORBING, particles=p, steps=60, laser=3.0e12)
    if case == "laser_sparse":
        # plasma only in a few cells left of the domain centre: the right-hand slabs start EMPTY and fill by migration
        nx, dx = 120, 2.0e-8
        p = pc.thermal_particles(nx, 12, dx, 0.4, seed=11, ne=5.0e26, cells=np.arange(45, 60))
        return dict(nx=nx, dx=dx, xmin=-1.0e-6, left=BDY_LASER, right=BDY_ABSORBING, particles=p, steps=60, laser=3.0e12)
    raise ValueError(case)


def laser_fns(amp):
    import math
    from opal_b200.constants import SPEED_OF_LIGHT as C
    omega = 2.0 * math.pi * C / 0.8e-6
    return (lambda t, x: amp * math.sin(omega * (t - x / C)), lambda t, x: 0.5 * amp * math.cos(omega * (t - x / C)))


def run_rank(rank, world, port, case, outdir, balanced):
    import torch.distributed as dist
    from opal_b200.abi import ELECTRON, OpalABI
    from opal_b200.slabs import GridDesign, SlabRing
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    oracle = OpalABI(ctypes.CDLL(str(ROOT / "oracle" / "libopal_oracle.so")), "opal_ref_")
    g = global_case(case)
    if balanced:
        cells = g["particles"]["cell"]
        dens = np.bincount(cells, minlength=g["nx"]).astype(float)
        ne = lambda x: dens[min(int(round((x - g["xmin"]) / g["dx"])), g["nx"] - 1)]
        design = GridDesign.balanced(world, g["nx"], g["xmin"], g["dx"], ne, left=g["left"], right=g["right"])
    else:
        design = GridDesign.unbalanced(world, g["nx"], g["xmin"], g["dx"], left=g["left"], right=g["right"])
    ring = SlabRing(oracle, design, rank, exchange="host", dist=dist)
    off = sum(design.nx[:rank])
    p = g["particles"]
    mine = (p["cell"] >= off) & (p["cell"] < off + design.nx[rank])
    ring.slab.upload(ELECTRON, p["cell"][mine] - off, p["x"][mine], p["ux"][mine], p["uy"][mine], p["uz"][mine], p["weight"][mine])
    t = 0.0
    laser = laser_fns(g["laser"]) if g["laser"] else None
    leavers = []
    for _ in range(g["steps"]):
        t = ring.step(t, laser=laser, species=(ELECTRON,))
        leavers.append(ring.slab.last_leavers(ELECTRON))
    d = ring.slab.download(ELECTRON)
    d["cell"] = d["cell"] + off
    e = ring.energy()
    np.savez(os.path.join(outdir, f"rank{rank}.npz"), fields=ring.slab.fields(), leavers=np.array(leavers), energy=np.array(e),
             nx=np.array(design.nx), left_size=ring.slab.left_size, **{"p_" + k: v for k, v in d.items()})
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    run_rank(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5], sys.argv[6] == "1")

"""Multi-GPU parity check, launched under torchrun (one rank per GPU):
the NCCL ring of CUDA slabs (in-library halo + migration exchange, fused run_steps) against ONE
CPU-oracle slab covering the whole domain.  Rank 0 prints 'MGPU OK ...' on success."""
import ctypes
import os
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def main():
    import torch
    import torch.distributed as dist
    import pic_cases as pc
    import mslab_worker as w
    from opal_b200 import lib
    from opal_b200.abi import ELECTRON, FLAG_LEAN_STATE, OpalABI, Slab
    from opal_b200.slabs import GridDesign, SlabRing

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    gpu = lib.load()
    ok = True
    cases = (("periodic", FLAG_LEAN_STATE, True), ("periodic", 0, True), ("laser", 0, False), ("laser_sparse", 0, False),
             ("laser_sparse", 0, True))
    for case, flags, fused in cases:
        g = w.global_case(case)
        design = GridDesign.unbalanced(world, g["nx"], g["xmin"], g["dx"], left=g["left"], right=g["right"])
        ring = SlabRing(gpu, design, rank, flags=flags, device=local, exchange="nccl", dist=dist)
        off = sum(design.nx[:rank])
        p = g["particles"]
        mine = (p["cell"] >= off) & (p["cell"] < off + design.nx[rank])
        ring.slab.upload(ELECTRON, p["cell"][mine] - off, p["x"][mine], p["ux"][mine], p["uy"][mine], p["uz"][mine], p["weight"][mine])
        laser = w.laser_fns(g["laser"]) if g["laser"] else None
        steps = g["steps"]
        if fused and laser:
            x2, dt = ring.slab.laser_x(), ring.slab.dt       # host-evaluated amplitudes, one per step (yee.rs:456-462; t += dt, main.rs:266)
            ts = np.cumsum(np.concatenate([[0.0], np.full(steps - 1, dt)]))
            ring.slab.run_steps(steps, 0.0, laser_y=[laser[0](t_, x2) for t_ in ts], laser_z=[laser[1](t_, x2) for t_ in ts])
        elif fused:
            ring.slab.run_steps(steps)
        else:
            t = 0.0
            for _ in range(steps):
                t = ring.step(t, laser=laser, species=(ELECTRON,))
        fields = ("cell", "x", "ux", "uy", "uz", "weight")
        d = ring.slab.download(ELECTRON, fields=fields)
        d["cell"] = d["cell"] + off
        f = ring.slab.fields()[ring.slab.left_size: ring.slab.left_size + design.nx[rank]]
        gathered = [None] * world
        dist.all_gather_object(gathered, (d, f))
        if rank == 0:
            oracle = OpalABI(ctypes.CDLL(str(ROOT / "oracle" / "libopal_oracle.so")), "opal_ref_")
            one = Slab(oracle, g["nx"], g["xmin"], g["dx"], None, g["left"], g["right"])
            pc.upload(one, ELECTRON, p)
            t = 0.0
            for _ in range(steps):
                t = pc.step_by_calls(one, t, laser=laser, species=(ELECTRON,))
            ref = one.download(ELECTRON, fields=fields)
            got = {k: np.concatenate([x[0][k] for x in gathered]) for k in fields}
            try:
                pc.compare_particles(got, ref, fields=("x", "ux", "uy", "uz"))
                covered = sum(design.nx)
                pc.compare_fields(np.concatenate([x[1] for x in gathered]), one.fields()[one.left_size: one.left_size + covered])
                print(f"MGPU OK case={case} flags={flags} fused={fused} world={world} particles={got['cell'].size}", flush=True)
            except AssertionError as e:
                ok = False
                print(f"MGPU FAIL case={case} flags={flags}: {e}", flush=True)
        ring.slab.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

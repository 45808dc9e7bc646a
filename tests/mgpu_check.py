"""Multi-GPU parity check, launched under torchrun (one rank per GPU): the NCCL ring of CUDA slabs (in-library halo +
migration exchange) against the in-process ring of CPU-oracle slabs of the SAME GridDesign (tests/ring_oracle.py).

Per case: after EVERY step the leaver counts (gone_left, gone_right) of GPU slab g equal those of oracle slab g bit for bit
(`Population::advance`'s return value, mod.rs:284 -- SURVEY 8e's contract); at the end cell keys per slab are bit-exact and
x, u and the fields agree to 1e-10.  Rank 0 prints one 'MGPU OK ...' line per case and exits 0 only if all passed.

Also importable: ring_case() / run_ring_parity() are what bench.py's `multi_gpu_parity` block runs."""
import os
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def ring_case(case):
    """global problem definitions (the small ones: mslab_worker.global_case)"""
    import mslab_worker as w
    import pic_cases as pc
    from opal_b200.abi import BDY_ABSORBING, BDY_INTERNAL, BDY_LASER
    if case == "periodic_1e6":
        # BASELINE config 5 in small: counter-streaming beams, warm enough that ~10 % of the particles change cell per step
        nx, npc, dx = 8192, 128, 500.0
        p = pc.two_stream_particles(nx, npc, dx, seed=21, thermal=0.25)
        return dict(nx=nx, dx=dx, xmin=0.0, left=BDY_INTERNAL, right=BDY_INTERNAL, particles=p, steps=100, laser=None, balanced=False)
    if case == "laser_balanced_1e6":
        # hole_boring.yaml in small: a plasma slab filling a quarter of the box, laser from the left, absorbing right edge;
        # GridDesign.balanced (main.rs:105-110 defaults to balance: true) gives unequal slabs
        nx, dx = 24000, 2.0e-8
        p = pc.thermal_particles(nx, 170, dx, 0.4, seed=23, ne=5.0e26, cells=np.arange(8000, 14000))
        return dict(nx=nx, dx=dx, xmin=-1.0e-4, left=BDY_LASER, right=BDY_ABSORBING, particles=p, steps=100, laser=3.0e12, balanced=True)
    g = w.global_case(case)
    g["balanced"] = False
    return g


def make_design(g, world):
    from opal_b200.slabs import GridDesign
    import ring_oracle
    if g["balanced"]:
        dens = ring_oracle.density_of(g["particles"], g["nx"])
        ne = lambda x: dens[min(int(round((x - g["xmin"]) / g["dx"])), g["nx"] - 1)]    # noqa: E731
        return GridDesign.balanced(world, g["nx"], g["xmin"], g["dx"], ne, left=g["left"], right=g["right"])
    return GridDesign.unbalanced(world, g["nx"], g["xmin"], g["dx"], left=g["left"], right=g["right"])


def run_ring_parity(gpu_abi, dist, rank, world, local, g, flags, fused, steps=None, per_step=True):
    """runs case `g` on the GPU ring (all ranks) and on the oracle ring (rank 0); returns a result dict on rank 0"""
    import mslab_worker as w
    import pic_cases as pc
    import ring_oracle
    from opal_b200.abi import ELECTRON
    from opal_b200.slabs import SlabRing
    design = make_design(g, world)
    steps = steps or g["steps"]
    ring = SlabRing(gpu_abi, design, rank, flags=flags, device=local, exchange="nccl", dist=dist)
    off = sum(design.nx[:rank])
    p = g["particles"]
    mine = (p["cell"] >= off) & (p["cell"] < off + design.nx[rank])
    ring.slab.upload(ELECTRON, p["cell"][mine] - off, p["x"][mine], p["ux"][mine], p["uy"][mine], p["uz"][mine], p["weight"][mine])
    laser = w.laser_fns(g["laser"]) if g["laser"] else None
    x2, dt = ring.slab.laser_x(), ring.slab.dt
    gone = []
    t = 0.0
    if fused and not per_step:
        ts = np.cumsum(np.concatenate([[0.0], np.full(steps - 1, dt)]))
        ly = [laser[0](t_, x2) for t_ in ts] if laser else None
        lz = [laser[1](t_, x2) for t_ in ts] if laser else None
        ring.slab.run_steps(steps, 0.0, laser_y=ly, laser_z=lz)
    else:
        for _ in range(steps):
            if fused:      # one fused step per call: the leaver counts of that step are readable afterwards
                ring.slab.run_steps(1, t, laser_y=[laser[0](t, x2)] if laser else None, laser_z=[laser[1](t, x2)] if laser else None)
                t += dt
            else:
                t = ring.step(t, laser=laser, species=(ELECTRON,))
            gone.append(ring.slab.last_leavers(ELECTRON))
    fields = ("cell", "x", "ux", "uy", "uz", "weight")
    d = ring.slab.download(ELECTRON, fields=fields)
    d["cell"] = d["cell"] + off
    f = ring.slab.fields()[ring.slab.left_size: ring.slab.left_size + design.nx[rank]]
    ring.slab.close()
    gathered = [None] * world
    dist.all_gather_object(gathered, (d, f, gone))
    if rank != 0:
        return None
    oracle = ring_oracle.OracleRing(ring_oracle.load_oracle(), design)
    oracle.upload(ELECTRON, p)
    ref_gone = []
    t = 0.0
    for _ in range(steps):
        t, gn = oracle.step(t, laser=laser, species=(ELECTRON,))
        ref_gone.append(gn[ELECTRON])
    res = {"slabs": list(design.nx), "particles": int(p["cell"].size), "steps": steps, "flags": flags, "fused": fused}
    try:
        if gone:
            got = np.array([x[2] for x in gathered]).transpose(1, 0, 2)          # [step][slab][side]
            want = np.array(ref_gone)
            res["leaver_counts_bit_exact"] = bool(np.array_equal(got, want))
            res["leavers_total"] = int(want.sum())
            assert res["leaver_counts_bit_exact"], f"per-step leaver counts differ first at step {int(np.argwhere((got != want).any(axis=(1, 2)))[0][0])}"
        ref_p, ref_f = oracle.download(ELECTRON, fields), oracle.interior_fields()
        worst_p, worst_f = {}, {}
        for r in range(world):                                                   # slab by slab: GPU g against oracle slab g
            wp = pc.compare_particles(gathered[r][0], ref_p[r], fields=("x", "ux", "uy", "uz"))
            wf = pc.compare_fields(gathered[r][1], ref_f[r])
            for k, v in wp.items():
                worst_p[k] = max(worst_p.get(k, 0.0), v)
            for k, v in wf.items():
                worst_f[k] = max(worst_f.get(k, 0.0), v)
        res.update({"cells_bit_exact": True, "max_rel_x_u": max(worst_p.values()), "max_rel_fields": max(worst_f.values()), "ok": True})
    except AssertionError as e:
        res.update({"ok": False, "error": str(e)[:300]})
    oracle.close()
    return res


def main():
    import torch
    import torch.distributed as dist
    from opal_b200 import lib
    from opal_b200.abi import FLAG_LEAN_STATE

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    gpu = lib.load()
    ok = True
    only = os.environ.get("MGPU_CASES")
    cases = (("periodic_1e6", FLAG_LEAN_STATE, True, True),          # the layout bench.py and SCALE run
             ("laser_balanced_1e6", FLAG_LEAN_STATE, True, True),    # unequal slabs (GridDesign.balanced), open edges
             ("periodic", FLAG_LEAN_STATE, True, True), ("periodic", 0, True, True), ("periodic", 0, True, False),
             ("laser", 0, False, True), ("laser_sparse", 0, False, True), ("laser_sparse", 0, True, True),
             # the neighbour exchange is unbounded (mod.rs:296-354 sends Vecs): outbox and inbox start at TWO records and grow
             ("periodic_tiny_boxes", FLAG_LEAN_STATE, True, True), ("periodic_tiny_boxes", 0, True, True))
    for case, flags, fused, per_step in cases:
        if only and case not in only.split(","):
            continue
        if case == "periodic_tiny_boxes":
            os.environ["OPAL_EXCHANGE_CAP"] = "2"
            case = "periodic"
        else:
            os.environ.pop("OPAL_EXCHANGE_CAP", None)
        res = run_ring_parity(gpu, dist, rank, world, local, ring_case(case), flags, fused, per_step=per_step)
        if rank == 0:
            ok = ok and res["ok"]
            print(f"MGPU {'OK' if res['ok'] else 'FAIL'} case={case} world={world} per_step={per_step} {res}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

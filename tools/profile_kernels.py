#!/usr/bin/env python3
"""Workloads for the per-kernel ncu captures kept under profiles/ (one phase per invocation, so that an `ncu -k regex:...`
run replays only what it is after).  No oracle, no torch: the product library through its C ABI.

  python tools/profile_kernels.py --phase lean_e   [--nx 100000]    k_push_lean_fast<.,.,ELECTRON>, k_pull      (1e8 electrons)
  python tools/profile_kernels.py --phase lean_i   [--nx 50000]     k_push_lean_fast<.,.,ION>, k_pull           (5e7 ions)
  python tools/profile_kernels.py --phase qed      [--nx 100000]    k_push<ELECTRON,FULL,.>, k_emit, k_scatter, k_photon_push (2e7 electrons)
  python tools/profile_kernels.py --phase two_stream                the whole two_stream.yaml-sized step (launch-bound)

ncu cannot replay the 54-GB launches of the bench size; per-particle figures from these sizes (same 1000 particles per
cell) are what profiles/push_traffic.json and bench.py's `roofline.traffic` scale up."""
import argparse
import pathlib
import sys
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

from opal_b200 import lib                                                   # noqa: E402
from opal_b200.abi import ELECTRON, FLAG_LEAN_STATE, ION, PHOTON, Slab      # noqa: E402

U_TH, NE, DX = 0.03128, 20.0, 500.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--phase", required=True, choices=["lean_e", "lean_i", "qed", "two_stream"])
    ap.add_argument("--nx", type=int, default=0)
    ap.add_argument("--npc", type=int, default=1000)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    abi = lib.load()
    t0 = time.perf_counter()
    if args.phase in ("lean_e", "lean_i"):
        nx = args.nx or (100000 if args.phase == "lean_e" else 50000)
        s = Slab(abi, nx, 0.0, DX, seed=12345, flags=FLAG_LEAN_STATE)
        if args.phase == "lean_e":
            s.generate(ELECTRON, args.npc, NE * DX / args.npc, u_th=(U_TH,) * 3, seed=12345)
        else:
            s.set_ion(6.0, 12.0)                                            # C6+ (examples/hole_boring.yaml:28-30)
            s.generate(ION, args.npc, NE * DX / args.npc / 6.0, u_th=(U_TH / 148.0,) * 3, seed=12345)
        f = s.fields()                                                      # a weak transverse wave so that E, B are not all zero
        k = 2.0 * np.pi * np.arange(f.shape[0]) / 64.0
        f[:, 6], f[:, 10] = 1.0e-3 * np.sin(k), 1.0e-3 * np.sin(k) / 2.997925e8
        s.set_fields(f)
        for _ in range(args.steps):
            s.run_steps(1)
        print(args.phase, "particles", s.count(ELECTRON) + s.count(ION), "energies", s.energy(), "launches", s.launch_count())
    elif args.phase == "qed":
        import bench_qed
        s = bench_qed.build(abi, args.nx or 100000, 200, 0.5)
        for k in range(args.steps):
            s.run_steps(1, t0=k * s.dt, current_deposition=False, photon_emission=True)
        print("qed electrons", s.count(ELECTRON), "photons", s.count(PHOTON), "launches", s.launch_count())
    else:
        s = Slab(abi, 1000, 0.0, DX, seed=0, flags=FLAG_LEAN_STATE)
        drift = 2.5e-24 / (9.109383e-31 * 2.997925e8)
        s.generate(ELECTRON, 100, NE * DX / 100, drift=(drift, 0.0, 0.0), two_stream=True, seed=1)
        s.run_steps(64)
        print("two_stream launches", s.launch_count(), "burst steps", s.burst_steps())
    s.close()
    print(f"{time.perf_counter() - t0:.1f} s")


if __name__ == "__main__":
    main()

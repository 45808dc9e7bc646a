#!/usr/bin/env python3
"""Throughput of the QED-on time step (BASELINE config 3 at scale): full particle state, photon emission on, current
deposition off as in examples/colliding_beams.yaml -- advance(e), advance(ph), emit_radiation, synchronize, field
advance through opal_run_steps (main.rs:238-267).  Electrons at gamma ~ 1000 meet a plane wave head-on (chi ~ --chi),
so a few per cent of them radiate every step; the photons stay in the periodic slab and are pushed with the rest.

  python tools/bench_qed.py [--nx 100000] [--epc 200] [--steps 20] [--cpu-nx 2000] [--out profiles/r01_qed_step.json]

Measurement harness, not product code: oracle/ is executed here only as the timed CPU baseline / checker beside the CUDA
path (the role bench.py's cpu_baseline leg has); nothing under opal_b200/ depends on it.
"""
import argparse
import ctypes
import json
import pathlib
import sys
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from opal_b200.abi import ELECTRON, PHOTON, OpalABI, Slab          # noqa: E402
from opal_b200.constants import CRITICAL_FIELD, SPEED_OF_LIGHT     # noqa: E402


def build(abi, nx, epc, chi, seed=5):
    dx, gamma0 = 1.0e-8, 1000.0
    s = Slab(abi, nx, 0.0, dx, seed=seed)
    e0 = chi * CRITICAL_FIELD / (2.0 * gamma0)
    f = s.fields()
    f[:, 6] = e0                         # Ey, Bz of a wave travelling towards +x (same columns as tools/bench_absorb.py)
    f[:, 10] = e0 / SPEED_OF_LIGHT
    s.set_fields(f)
    rng = np.random.default_rng(seed)
    ne = nx * epc
    s.upload(ELECTRON, np.repeat(np.arange(nx), epc), rng.random(ne), -gamma0 * (1.0 + 0.01 * rng.standard_normal(ne)),
             np.zeros(ne), np.zeros(ne), np.full(ne, 1.0e10), tau=rng.exponential(size=ne))
    return s


def timed(s, steps, warmup):
    s.run_steps(warmup, current_deposition=False, photon_emission=True)
    n0 = s.count(PHOTON)
    t0 = time.perf_counter()
    s.run_steps(steps, t0=warmup * s.dt, current_deposition=False, photon_emission=True)   # returns after the status read-back
    dt = (time.perf_counter() - t0) / steps
    n1 = s.count(PHOTON)
    return dt, n0, n1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=100000)
    ap.add_argument("--epc", type=int, default=200)
    ap.add_argument("--chi", type=float, default=0.5)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--cpu-nx", type=int, default=2000)
    ap.add_argument("--cpu-steps", type=int, default=5)
    ap.add_argument("--out", default=str(ROOT / "profiles" / "r01_qed_step.json"))
    args = ap.parse_args()
    from opal_b200 import lib
    gpu = lib.load()
    s = build(gpu, args.nx, args.epc, args.chi)
    ne = s.count(ELECTRON)
    dt_gpu, n0, n1 = timed(s, args.steps, args.warmup)
    en = s.energy()
    s.close()
    line = {"what": "QED-on step (full state, emission on, deposition off): electron-steps per second; photons are pushed on top",
            "chi": args.chi,
            "gpu": {"electron_steps_per_s": ne / dt_gpu, "ms_per_step": 1e3 * dt_gpu, "electrons": ne, "nx": args.nx,
                    "photons_before": n0, "photons_after": n1, "photons_emitted_per_step": (n1 - n0) / args.steps,
                    "particle_steps_per_s_incl_photons": (ne + 0.5 * (n0 + n1)) / dt_gpu, "energies_J": list(en)}}
    if args.cpu_nx > 0:
        oracle = OpalABI(ctypes.CDLL(str(next((ROOT / "oracle").glob("libopal*.so")))), "opal_ref_")
        o = build(oracle, args.cpu_nx, args.epc, args.chi)
        nec = o.count(ELECTRON)
        dt_cpu, m0, m1 = timed(o, args.cpu_steps, 1)
        o.close()
        line["cpu_oracle"] = {"electron_steps_per_s": nec / dt_cpu, "ms_per_step": 1e3 * dt_cpu, "electrons": nec, "nx": args.cpu_nx,
                              "photons_emitted_per_step": (m1 - m0) / args.cpu_steps, "kind": "port (OpenMP, all host cores)"}
        line["ratio"] = line["gpu"]["electron_steps_per_s"] / line["cpu_oracle"]["electron_steps_per_s"]
    print(json.dumps(line))
    pathlib.Path(args.out).write_text(json.dumps(line, indent=1) + "\n")


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""Throughput of absorb (particle/interactions.rs:145) in photon-electron pairs per second: the CUDA library on the GPU and
the oracle on one host core, on the same kind of population (every cell holds `epc` electrons and `ppc` photons, fields
of a plane wave at chi ~ 1, optical depths too large for any event so that every pair of a cell is evaluated).

  python tools/bench_absorb.py [--nx 4096] [--epc 64] [--ppc 64] [--cpu-nx 128] [--out profiles/r01_absorb.json]

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


def build(abi, nx, epc, ppc, seed=3):
    dx, gamma0 = 1.0e-8, 1000.0
    s = Slab(abi, nx, 0.0, dx, 0.1 * dx / SPEED_OF_LIGHT, seed=seed)
    e0 = CRITICAL_FIELD / (2.0 * gamma0)
    f = s.fields()
    f[:, 6] = e0
    f[:, 10] = e0 / SPEED_OF_LIGHT
    s.set_fields(f)
    rng = np.random.default_rng(seed)
    ne, nph = nx * epc, nx * ppc
    s.upload(ELECTRON, np.repeat(np.arange(nx), epc), 0.3 + 0.4 * rng.random(ne), -gamma0 * (1.0 + 0.2 * rng.random(ne)),
             3.0 * rng.standard_normal(ne), 3.0 * rng.standard_normal(ne), np.full(ne, 1.0e24))
    k0 = gamma0 * 10.0 ** rng.uniform(-3.0, 0.2, nph)
    th = 2.0e-3 * rng.standard_normal(nph)
    s.upload(PHOTON, np.repeat(np.arange(nx), ppc), 0.3 + 0.4 * rng.random(nph), -k0 * np.cos(th), k0 * np.sin(th), np.zeros(nph), np.ones(nph))
    s.advance(ELECTRON)
    s.advance(PHOTON)
    s.set_photon_pair_depths(np.full(nph, 1.0e3), np.full(nph, 1.0e3))
    return s


def timed(s, calls):
    s.absorb(0.0)                        # warm-up (allocates the event arrays)
    t0 = time.perf_counter()
    for _ in range(calls):
        s.absorb(0.0)                    # returns after the status read-back: the call is complete
    return (time.perf_counter() - t0) / calls


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=4096)
    ap.add_argument("--epc", type=int, default=64)
    ap.add_argument("--ppc", type=int, default=64)
    ap.add_argument("--cpu-nx", type=int, default=128)
    ap.add_argument("--calls", type=int, default=10)
    ap.add_argument("--out", default=str(ROOT / "profiles" / "r01_absorb.json"))
    args = ap.parse_args()
    from opal_b200 import lib
    gpu = lib.load()
    # the last occupied cell never interacts (interactions.rs:216-222): nx - 1 cells of pairs
    s = build(gpu, args.nx, args.epc, args.ppc)
    dt_gpu = timed(s, args.calls)
    pairs_gpu = (args.nx - 1) * args.epc * args.ppc
    spent = 1.0e3 - s.photon_pair_depths()[0]
    s.close()
    oracle = OpalABI(ctypes.CDLL(str(next((ROOT / "oracle").glob("libopal*.so")))), "opal_ref_")
    o = build(oracle, args.cpu_nx, args.epc, args.ppc)
    dt_cpu = timed(o, 1)
    pairs_cpu = (args.cpu_nx - 1) * args.epc * args.ppc
    o.close()
    line = {"what": "absorb: photon-electron pairs evaluated per second (both cross sections per pair)",
            "gpu": {"pairs_per_s": pairs_gpu / dt_gpu, "ms_per_call": 1e3 * dt_gpu, "pairs_per_call": pairs_gpu, "nx": args.nx,
                    "electrons_per_cell": args.epc, "photons_per_cell": args.ppc, "photons_with_pairs": int((spent > 0).sum())},
            "cpu_oracle_1_core": {"pairs_per_s": pairs_cpu / dt_cpu, "ms_per_call": 1e3 * dt_cpu, "pairs_per_call": pairs_cpu, "nx": args.cpu_nx}}
    line["ratio"] = line["gpu"]["pairs_per_s"] / line["cpu_oracle_1_core"]["pairs_per_s"]
    print(json.dumps(line))
    pathlib.Path(args.out).write_text(json.dumps(line, indent=1) + "\n")


if __name__ == "__main__":
    main()

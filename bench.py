#!/usr/bin/env python3
"""bench.py -- particle-steps/s of the opal PIC hot path (push + deposit + field) on B200.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle port)

Workload (BASELINE.json config 4, per GPU): synthetic thermal plasma, nx = 1e6 cells, npc = 1000
-> 1e9 electron macroparticles, dx = 500, ne = 20, u = 0.03128 * N(0,1) on all three components,
periodic, current deposition on, QED off, lean (x, u, w) f64 SoA layout; a step is one iteration
of src/main.rs:238-267.  For N > 1 every rank owns such a slab of a ring (weak scaling): NCCL
halo + particle migration, no other collective.

`value`  : whole-job particle-steps/s, device-timed (CUDA events on the library's stream), state
           resident in HBM, max over ranks.
`e2e`    : the same metric through the per-step public call a driver makes: each step passes the
           host-evaluated laser amplitudes (host -> device) and reads the four conservation
           energies + status back (device -> host); wall clock.
`roofline`: 80 algorithmic bytes per particle-step x particles per launch / mean
           duration of the fused gather+push+deposit kernel, against MEASURED_PEAKS.json hbm_gbs.
`cpu_baseline`: the CPU oracle (C++ restatement of opal's Rayon path; the Rust binary cannot be
           built here) on a bounded sample, all host threads.
"""
import argparse
import ctypes
import json
import os
import pathlib
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

B_ALG = 80.0          # bytes per particle-step: read i32 cell + x + u[3] + w = 44, write 36
U_TH = 0.03128
NE = 20.0
DX = 500.0


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """samples nvidia-smi clocks / throttle reasons during the timed region"""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([v.strip() for v in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 7 for k in range(4) if r[3 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def load_oracle():
    """cpu_baseline / --impl reference only: the CPU oracle (never on the product path)."""
    so = ROOT / "oracle" / "libopal_oracle.so"
    if not so.exists():
        subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    from opal_b200.abi import OpalABI
    cdll = ctypes.CDLL(str(so))
    return cdll, OpalABI(cdll, "opal_ref_")


def cpu_reference_run(nx, npc, steps, warmup, threads=None):
    """time the oracle ("reference-shaped": AoS structs, full sort every step, per-thread private
    deposit grids + serial reduce, serial field loops) on a bounded sample of the bench workload"""
    from opal_b200.abi import ELECTRON, Slab
    cdll, oracle = load_oracle()
    cdll.opal_ref_max_threads.restype = ctypes.c_int
    # all host cores this process may run on: torchrun exports OMP_NUM_THREADS=1 for N > 1, which would
    # shrink omp_get_max_threads() to one thread although rank 0 alone runs the reference arm
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    threads = threads or int(os.environ.get("OPAL_CPU_THREADS", max(avail, cdll.opal_ref_max_threads())))
    cdll.opal_ref_set_threads(threads)
    rng = np.random.default_rng(12345)
    n = nx * npc
    s = Slab(oracle, nx, 0.0, DX)
    s.upload(ELECTRON, np.repeat(np.arange(nx, dtype=np.int64), npc), rng.random(n), U_TH * rng.standard_normal(n),
             U_TH * rng.standard_normal(n), U_TH * rng.standard_normal(n), np.full(n, NE * DX / npc))
    s.run_steps(warmup)
    t0 = time.perf_counter()
    s.run_steps(steps, warmup * s.dt)
    dt = time.perf_counter() - t0
    s.close()
    return n * steps / dt, dt, threads, n


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nx, npc = args.cpu_nx, args.npc
    rate, secs, threads, n = cpu_reference_run(nx, npc, args.steps, args.warmup)
    sample = (f"{n:.3g} electrons / {nx} cells (npc {npc}), {args.steps} steps after {args.warmup} warm-up: a bounded sample of the "
              f"1e9-particle / 1e6-cell workload, same density, temperature and ppc")
    line = {
        "impl": "reference", "metric": "particle-steps/sec (push+deposit+field)", "value": rate, "unit": "particle-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "synthetic thermal plasma (BASELINE config 4), CPU-bounded sample", "particles": n, "cells": nx,
                   "npc": npc, "reference": "C++ restatement of opal's mpirun+Rayon CPU path (oracle/); the Rust binary cannot be built on this box"},
        "cpu_baseline": {"value": rate, "unit": "particle-steps/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "particle-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nx", type=int, default=1_000_000, help="cells per GPU")
    ap.add_argument("--npc", type=int, default=1000)
    ap.add_argument("--workload", default="thermal", choices=["thermal", "two_stream"])
    ap.add_argument("--cpu-nx", type=int, default=40000, help="cells of the bounded CPU sample (x npc particles; about 10-30 s of host work)")
    ap.add_argument("--cpu-steps", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--full-state", action="store_true", help="keep all reference per-particle fields (152 B/particle-step)")
    args = ap.parse_args()

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import torch.distributed as dist
    from opal_b200 import lib
    from opal_b200.abi import ELECTRON, FLAG_LEAN_STATE, Slab

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    abi = lib.load()

    nx, npc = args.nx, args.npc
    n_local = nx * npc
    flags = 0 if args.full_state else FLAG_LEAN_STATE
    slab = Slab(abi, nx, rank * nx * DX, DX, rank=rank, nranks=world, seed=12345, flags=flags, device=local)
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = ctypes.create_string_buffer(128)
            assert abi.comm_unique_id(buf) == 0
            uid.copy_(torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        slab.comm_init(bytes(uid.cpu().numpy().tobytes()))
    if args.workload == "thermal":
        slab.generate(ELECTRON, npc, NE * DX / npc, u_th=(U_TH,) * 3, global_cell_offset=rank * nx, seed=12345)
    else:
        from opal_b200.constants import ELECTRON_MASS, SPEED_OF_LIGHT
        drift = 2.5e-24 / (ELECTRON_MASS * SPEED_OF_LIGHT)
        slab.generate(ELECTRON, npc, NE * DX / npc, drift=(drift, 0.0, 0.0), two_stream=True, global_cell_offset=rank * nx, seed=12345)

    stream = torch.cuda.ExternalStream(slab.stream(), device=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-timed arm ---------------------------------------------------------------
    slab.run_steps(args.warmup)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = slab.launch_count()
    slab.timing(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    slab.run_steps(args.steps, args.warmup * slab.dt)
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    kt = slab.kernel_times()
    slab.timing(False)
    launches = slab.launch_count() - l0

    # ---- end-to-end arm: per-step public call with host laser scalars in, energies out ---------
    barrier()
    e2e_steps = args.steps
    t0 = time.perf_counter()
    t = (args.warmup + args.steps) * slab.dt
    zeros = np.zeros(1)
    energies = None
    for _ in range(e2e_steps):
        slab.run_steps(1, t, laser_y=zeros, laser_z=zeros)     # laser amplitudes evaluated on the host (yee.rs:456-462)
        energies = slab.energy()                                # write_energies inputs (main.rs:23-42)
        t += slab.dt
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None      # sampled over both timed arms: the GPU is under the same load in each

    tt = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = tt.tolist()
    total_particles = n_local * world
    value = total_particles * args.steps / (ms_max * 1e-3)
    e2e_value = total_particles * e2e_steps / (e2e_ms_max * 1e-3)

    if rank == 0:
        peak, peak_src = peaks()
        push_ms, push_n = kt["push"]
        achieved = (B_ALG * n_local / (push_ms / push_n * 1e-3)) / 1e9 if push_n else None
        traffic = None
        tfile = ROOT / "profiles" / "push_traffic.json"
        if tfile.exists() and not args.full_state:
            try:
                traffic = json.loads(tfile.read_text())["dram_bytes_per_particle"] * n_local   # ncu dram bytes per launch
            except Exception:
                pass
        line = {
            "metric": "particle-steps/sec (push+deposit+field)", "value": value, "unit": "particle-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"synthetic {args.workload} plasma (BASELINE config {'4' if args.workload == 'thermal' else '5'}), "
                                   f"{n_local:.3g} electrons / {nx} cells per GPU, periodic ring of x-slabs",
                       "particles_per_gpu": n_local, "cells_per_gpu": nx, "npc": npc, "layout": "full" if args.full_state else "lean x,u,w f64 SoA",
                       "l2": "inputs (>= 40 GB per step) far larger than the 126 MB L2; no flush needed",
                       "parallelism": f"x-slab ring x{world}"},
            "hbm_gbs_algorithmic": value / world * B_ALG / 1e9,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if achieved else None,
                         "traffic": traffic, "peak_source": peak_src, "kernel": "k_push<electron, lean, fused deposit>",
                         "algorithmic_bytes_per_particle_step": B_ALG,
                         "kernel_ms": {k: (v[0] / v[1] if v[1] else None) for k, v in kt.items()},
                         "kernel_share_of_step": {k: v[0] / ms for k, v in kt.items()}},
            "e2e": {"value": e2e_value, "unit": "particle-steps/s", "h2d_bytes_per_step": 16, "d2h_bytes_per_step": 32 + 44,
                    "note": "per step: 2 host-evaluated laser amplitudes in, 4 energies + status block out; particle state stays resident"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "energies_J": energies,
        }
        if not args.no_cpu_baseline and world == 1:   # the CPU leg is timed on rank 0 at N = 1 only
            cnx = min(args.cpu_nx, nx)
            rate, secs, threads, n = cpu_reference_run(cnx, npc, args.cpu_steps, 1)
            line["cpu_baseline"] = {"value": rate, "unit": "particle-steps/s", "cores": threads, "kind": "port",
                                    "sample": f"{n:.3g} electrons / {cnx} cells (same ppc, density, temperature), {args.cpu_steps} steps, {secs:.1f} s; "
                                              "C++ restatement of opal's mpirun+Rayon CPU path -- the Rust binary is unavailable on this box"}
        print(json.dumps(line), flush=True)
    slab.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""bench.py -- particle-steps/s of the opal PIC hot path (push + deposit + field) on B200.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle port)

Workload (BASELINE.json config 4, per GPU): synthetic thermal plasma, nx = 1e6 cells, npc = 1000
-> 1e9 electron macroparticles, dx = 500, ne = 20, u = 0.03128 * N(0,1) on all three components,
periodic, current deposition on, QED off, lean (x, u, w) f64 SoA layout; a step is one iteration
of src/main.rs:238-267.  For N > 1 every rank owns such a slab of a ring (weak scaling): NCCL
halo + particle migration, no other collective.

Protocol: W warm-up steps, then the block of EXACTLY K timed steps is repeated
--reps times (default 5) inside this one invocation and the MEDIAN block is reported (`ms_per_step`,
`value`; all blocks in `reps_ms_per_step`); the clock sampler (NVML, ~50 Hz) runs across all blocks.

`value`  : whole-job particle-steps/s, device-timed (CUDA events on the library's stream), state
           resident in HBM, max over ranks.
`e2e`    : the same metric through the per-step public call a driver makes: each step passes the
           host-evaluated laser amplitudes (host -> device) and reads the four conservation
           energies + status back (device -> host); wall clock.
`roofline`: 80 algorithmic bytes per particle-step x particles per launch / mean
           duration of the fused gather+push+deposit kernel, against MEASURED_PEAKS.json hbm_gbs.
`cpu_baseline`: the CPU oracle (C++ restatement of opal's Rayon path; the Rust binary cannot be
           built here) on a bounded sample, all host threads.
`parity` : (N = 1) the very arrays of the CPU leg (4e7 electrons) through the lean CUDA path for the
           same steps: cell keys bit-exact, x/u/fields norm-wise relative error (tolerance 1e-10).
`multi_gpu_parity`: (N > 1) before the timed region, a 1e6-particle lean ring on the N GPUs against
           the in-process ring of N oracle slabs on rank 0: per-step leaver counts bit-exact.
`config5`: the same protocol on BASELINE config 5 (two-stream drift of examples/two_stream.yaml:18-20,
           1e9 electrons per GPU) -- the weak-scaling deck, at every N including 1.
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

    def _nvml_loop(self):
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(self.index)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self._stop_evt.is_set():
            sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            try:
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
            except Exception:
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            pw = nv.nvmlDeviceGetPowerUsage(h) / 1e3
            self.rows.append([str(sm), str(mx), f"{pw:.1f}"] + ["Active" if r & bits[k] else "Not Active" for k in
                                                                 ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")])
            self._stop_evt.wait(0.02)

    def run(self):
        try:
            self._nvml_loop()
            return
        except Exception:
            pass                                  # no NVML binding: fall back to nvidia-smi (slower sampling)
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
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows), "power_w_median": float(np.median(pw)) if pw else None}


def load_oracle():
    """cpu_baseline / --impl reference only: the CPU oracle (never on the product path)."""
    so = ROOT / "oracle" / "libopal_oracle.so"
    if not so.exists():
        subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    from opal_b200.abi import OpalABI
    cdll = ctypes.CDLL(str(so))
    return cdll, OpalABI(cdll, "opal_ref_")


def cpu_reference_run(nx, npc, steps, warmup, threads=None, keep_state=False):
    """time the oracle ("reference-shaped": AoS structs, full sort every step, per-thread private
    deposit grids + serial reduce, serial field loops) on a bounded sample of the bench workload.
    keep_state: also return the initial arrays and the oracle's final particles / fields (parity leg)."""
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
    # every particle carries a unique weight (1e-9 relative steps): its identity in the parity comparison
    init = dict(cell=np.repeat(np.arange(nx, dtype=np.int64), npc), x=rng.random(n), ux=U_TH * rng.standard_normal(n),
                uy=U_TH * rng.standard_normal(n), uz=U_TH * rng.standard_normal(n), weight=(NE * DX / npc) * (1.0 + 1.0e-9 * np.arange(n)))
    s.upload(ELECTRON, init["cell"], init["x"], init["ux"], init["uy"], init["uz"], init["weight"])
    s.run_steps(warmup)
    t0 = time.perf_counter()
    s.run_steps(steps, warmup * s.dt)
    dt = time.perf_counter() - t0
    state = None
    if keep_state:
        state = {"init": init, "particles": s.download(ELECTRON, fields=("cell", "x", "ux", "uy", "uz", "weight")), "fields": s.fields(),
                 "energy": s.energy(), "steps": warmup + steps}
    s.close()
    return n * steps / dt, dt, threads, n, state


def rel_inf(a, b):
    """norm-wise relative error ||a - b||_inf / ||b||_inf (SURVEY 8c's parity criterion)"""
    scale = float(np.max(np.abs(b))) if b.size else 0.0
    diff = float(np.max(np.abs(a - b))) if b.size else 0.0
    return 0.0 if diff == 0.0 else diff / (scale if scale > 0.0 else 1.0)


def gpu_parity_leg(abi, state, nx, device):
    """the CPU leg's arrays through the lean CUDA path for the same steps, compared with the oracle's result"""
    from opal_b200.abi import ELECTRON, FLAG_LEAN_STATE, Slab
    fields = ("cell", "x", "ux", "uy", "uz", "weight")
    g = Slab(abi, nx, 0.0, DX, flags=FLAG_LEAN_STATE, device=device)
    i = state["init"]
    g.upload(ELECTRON, i["cell"], i["x"], i["ux"], i["uy"], i["uz"], i["weight"])
    g.run_steps(state["steps"])
    got, gf, ge = g.download(ELECTRON, fields=fields), g.fields(), g.energy()
    g.close()
    ref, rf, re_ = state["particles"], state["fields"], state["energy"]

    def canon(d):
        order = np.lexsort((d["weight"], d["cell"]))
        return {k: v[order] for k, v in d.items()}
    out = {"n": int(i["cell"].size), "steps": state["steps"], "tolerance": 1.0e-10}
    if got["cell"].size != ref["cell"].size:
        out.update({"cells_bit_exact": False, "ok": False, "error": "particle counts differ"})
        return out
    a, b = canon(got), canon(ref)
    out["cells_bit_exact"] = bool(np.array_equal(a["cell"], b["cell"]) and np.array_equal(a["weight"], b["weight"]))
    out["max_rel_x_u"] = max(rel_inf(a[k], b[k]) for k in ("x", "ux", "uy", "uz"))
    names = ["rho", "jx", "jy", "jz", "Ex", "Ey", "Ez", "Bx", "By", "Bz"]
    out["max_rel_fields"] = max(rel_inf(gf[:, k + 1], rf[:, k + 1]) for k in range(len(names)))
    out["max_rel_energies"] = max(abs(x - y) / abs(y) for x, y in zip(ge[:2], re_[:2]) if y)
    out["ok"] = bool(out["cells_bit_exact"] and out["max_rel_x_u"] <= 1e-10 and out["max_rel_fields"] <= 1e-10)
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nx, npc = args.cpu_nx, args.npc
    rate, secs, threads, n, _ = cpu_reference_run(nx, npc, args.steps, args.warmup)
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


def timed_blocks(slab, stream, barrier, steps, reps, t0):
    """`reps` blocks of exactly `steps` fused steps each, device-timed on the library's stream; returns ms per block"""
    import torch
    out = []
    t = t0
    for _ in range(reps):
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        slab.run_steps(steps, t)
        ev1.record(stream)
        barrier()
        out.append(ev0.elapsed_time(ev1))
        t += steps * slab.dt
    return out, t


def e2e_blocks(slab, barrier, steps, reps, t0):
    """the same blocks through the per-step public call: laser amplitudes in (host -> device), energies + status out"""
    out, t, energies = [], t0, None
    zeros = np.zeros(1)
    for _ in range(reps):
        barrier()
        w0 = time.perf_counter()
        for _ in range(steps):
            slab.run_steps(1, t, laser_y=zeros, laser_z=zeros)     # laser amplitudes evaluated on the host (yee.rs:456-462)
            energies = slab.energy()                                # write_energies inputs (main.rs:23-42)
            t += slab.dt
        barrier()
        out.append((time.perf_counter() - w0) * 1e3)
    return out, t, energies


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--reps", type=int, default=5, help="timed blocks of --steps steps; the median block is reported")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nx", type=int, default=1_000_000, help="cells per GPU")
    ap.add_argument("--npc", type=int, default=1000)
    ap.add_argument("--workload", default="thermal", choices=["thermal", "two_stream"])
    ap.add_argument("--cpu-nx", type=int, default=40000, help="cells of the bounded CPU sample (x npc particles; about 10-30 s of host work)")
    ap.add_argument("--cpu-steps", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity / multi_gpu_parity blocks")
    ap.add_argument("--no-config5", action="store_true", help="skip the second workload (BASELINE config 5)")
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

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- N > 1: correctness of the ring BEFORE anything is timed -----------------------------------
    mgpu_parity = None
    if world > 1 and not args.no_parity:
        sys.path.insert(0, str(ROOT / "tests"))
        import mgpu_check
        mgpu_parity = mgpu_check.run_ring_parity(abi, dist, rank, world, local, mgpu_check.ring_case("periodic_1e6"), FLAG_LEAN_STATE,
                                                 fused=True, steps=50, per_step=True)
        barrier()

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

    from opal_b200.constants import ELECTRON_MASS, SPEED_OF_LIGHT
    drift = 2.5e-24 / (ELECTRON_MASS * SPEED_OF_LIGHT)          # examples/two_stream.yaml:27

    def generate(workload):
        if workload == "thermal":
            slab.generate(ELECTRON, npc, NE * DX / npc, u_th=(U_TH,) * 3, global_cell_offset=rank * nx, seed=12345)
        else:
            slab.generate(ELECTRON, npc, NE * DX / npc, drift=(drift, 0.0, 0.0), two_stream=True, global_cell_offset=rank * nx, seed=12345)

    stream = torch.cuda.ExternalStream(slab.stream(), device=torch.device("cuda", local))
    total_particles = n_local * world

    def reduce_max(v):
        tt = torch.tensor(v, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return tt.tolist()

    def measure(workload, with_e2e):
        generate(workload)
        slab.run_steps(args.warmup)
        l0 = slab.launch_count()
        slab.timing(True)
        blocks, t = timed_blocks(slab, stream, barrier, args.steps, args.reps, args.warmup * slab.dt)
        kt = slab.kernel_times()
        slab.timing(False)
        launches = slab.launch_count() - l0
        e2e, energies = None, None
        if with_e2e:
            e2e, t, energies = e2e_blocks(slab, barrier, args.steps, args.reps, t)
            e2e = reduce_max(e2e)
        blocks = reduce_max(blocks)                     # per block: the slowest rank
        return {"blocks_ms": blocks, "ms": float(np.median(blocks)), "kt": kt, "launches": launches,
                "e2e_ms": float(np.median(e2e)) if e2e else None, "energies": energies}

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    main_m = measure(args.workload, True)
    clocks = sampler.stop() if rank == 0 else None       # sampled across every timed block of both arms
    other = "two_stream" if args.workload == "thermal" else "thermal"
    other_m = None if (args.no_config5 or args.full_state) else measure(other, False)
    slab.close()

    if rank == 0:
        ms = main_m["ms"]
        kt = main_m["kt"]
        value = total_particles * args.steps / (ms * 1e-3)
        e2e_value = total_particles * args.steps / (main_m["e2e_ms"] * 1e-3)
        peak, peak_src = peaks()
        push_ms, push_n = kt["push"]
        achieved = (B_ALG * n_local / (push_ms / push_n * 1e-3)) / 1e9 if push_n else None
        traffic, traffic_src = None, None
        tfile = ROOT / "profiles" / "push_traffic.json"
        if tfile.exists() and not args.full_state:
            try:
                tj = json.loads(tfile.read_text())
                traffic = tj["dram_bytes_per_particle"] * n_local   # ncu dram bytes read + written per launch
                traffic_src = tj.get("source")
            except Exception:
                pass
        total_timed_ms = float(sum(main_m["blocks_ms"]))

        def cfg(workload):
            return (f"synthetic {workload} plasma (BASELINE config {'4' if workload == 'thermal' else '5'}), "
                    f"{n_local:.3g} electrons / {nx} cells per GPU, periodic ring of x-slabs")
        line = {
            "metric": "particle-steps/sec (push+deposit+field)", "value": value, "unit": "particle-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": cfg(args.workload),
                       "particles_per_gpu": n_local, "cells_per_gpu": nx, "npc": npc, "layout": "full" if args.full_state else "lean x,u,w f64 SoA",
                       "l2": "inputs (>= 40 GB per step) far larger than the 126 MB L2; no flush needed",
                       "parallelism": f"x-slab ring x{world}",
                       "protocol": f"{args.warmup} warm-up steps, then {args.reps} blocks of {args.steps} timed steps; median block reported"},
            "reps": args.reps, "reps_ms_per_step": [b / args.steps for b in main_m["blocks_ms"]],
            "hbm_gbs_algorithmic": value / world * B_ALG / 1e9,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if achieved else None,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "kernel": "k_push_lean_fast<electron> (fused gather + Vay push + deposit)",
                         "algorithmic_bytes_per_particle_step": B_ALG,
                         "kernel_ms": {k: (v[0] / v[1] if v[1] else None) for k, v in kt.items()},
                         "kernel_share_of_step": {k: v[0] / total_timed_ms for k, v in kt.items()}},
            "e2e": {"value": e2e_value, "unit": "particle-steps/s", "h2d_bytes_per_step": 16, "d2h_bytes_per_step": 32 + 44,
                    "note": "per step: 2 host-evaluated laser amplitudes in, 4 energies + status block out; particle state stays resident"},
            "gpu_launches": int(main_m["launches"]),
            "clocks": clocks,
            "energies_J": main_m["energies"],
        }
        if other_m:
            key = "config5" if other == "two_stream" else "config4"
            line[key] = {"workload": cfg(other), "value": total_particles * args.steps / (other_m["ms"] * 1e-3), "unit": "particle-steps/s",
                         "ms_per_step": other_m["ms"] / args.steps, "reps_ms_per_step": [b / args.steps for b in other_m["blocks_ms"]],
                         "kernel_ms": {k: (v[0] / v[1] if v[1] else None) for k, v in other_m["kt"].items()}}
        if mgpu_parity is not None:
            line["multi_gpu_parity"] = mgpu_parity
        if world == 1 and not (args.no_cpu_baseline and args.no_parity):   # the CPU leg is timed on rank 0 at N = 1 only
            cnx = min(args.cpu_nx, nx)
            rate, secs, threads, n, state = cpu_reference_run(cnx, npc, args.cpu_steps, 1, keep_state=not args.no_parity)
            if not args.no_cpu_baseline:
                line["cpu_baseline"] = {"value": rate, "unit": "particle-steps/s", "cores": threads, "kind": "port",
                                        "sample": f"{n:.3g} electrons / {cnx} cells (same ppc, density, temperature), {args.cpu_steps} steps, {secs:.1f} s; "
                                                  "C++ restatement of opal's mpirun+Rayon CPU path -- the Rust binary is unavailable on this box"}
            if state is not None:
                line["parity"] = gpu_parity_leg(abi, state, cnx, local)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

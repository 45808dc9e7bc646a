#!/usr/bin/env python3
"""Runs the three BASELINE example configurations (two_stream, hole_boring, colliding_beams: SURVEY section 8 size table)
end to end through the C++ driver on the GPU, to completion, and the same deck through the oracle build of the driver up
to the first output for a full-size parity check.  Prints one JSON line per deck (also collected in --out).

The decks are the reference's own files, kept byte for byte under examples/ (/root/reference does not travel to the GPU box).
--qed-seeds N additionally runs colliding_beams.yaml at full size N times per side (GPU driver / oracle driver, --seed 1..N)
and compares photon yield, spectra and the electrons' energy loss at the last output (tests/qed_deck_stats.py).

Measurement harness, not product code: oracle/ is executed here only as the timed CPU baseline / checker beside the CUDA
path (the role bench.py's cpu_baseline leg has); nothing under opal_b200/ depends on it."""
import argparse
import json
import os
import pathlib
import subprocess
import sys
import time

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))

EXAMPLES = ROOT / "examples"      # the reference's examples/*.yaml, byte for byte (tests/test_examples_verbatim.py)


def deck_text(name):
    return (EXAMPLES / f"{name}.yaml").read_text()


DECKS = {"two_stream": lambda: deck_text("two_stream"), "hole_boring": lambda: deck_text("hole_boring"),
         "colliding_beams": lambda: deck_text("colliding_beams")}


def energies(path):
    return {k: float(v) for k, v in (line.split() for line in pathlib.Path(path).read_text().splitlines())}


def run(driver, workdir, extra=(), env=None, timeout=3000):
    t0 = time.perf_counter()
    p = subprocess.run([str(driver), str(workdir / "in.yaml"), *extra], capture_output=True, text=True, timeout=timeout, env=env)
    wall = time.perf_counter() - t0
    if p.returncode != 0:
        raise RuntimeError(f"{driver} failed:\n{p.stdout[-1500:]}\n{p.stderr[-1500:]}")
    rate = None
    for line in p.stdout.splitlines():
        if line.startswith("time loop:"):
            rate = float(line.split(",")[1].split()[0])
    return wall, rate, p.stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--work", default="/tmp/opal_examples")
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "examples.jsonl"))
    ap.add_argument("--decks", default="two_stream,hole_boring,colliding_beams")
    ap.add_argument("--ref-outputs", type=int, default=1, help="outputs the oracle build runs to (0: skip the oracle arm)")
    ap.add_argument("--ref-threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--write-only", action="store_true", help="only write the decks")
    ap.add_argument("--qed-seeds", type=int, default=0, help="full-size colliding_beams.yaml, this many seeds per side, statistics at the last output")
    args = ap.parse_args()
    work = pathlib.Path(args.work)
    lines = []
    for name in args.decks.split(","):
        for arm in ("gpu", "ref"):
            (work / name / arm).mkdir(parents=True, exist_ok=True)
            (work / name / arm / "in.yaml").write_text(DECKS[name]())
        if args.write_only:
            continue
        try:
            gpu_wall, gpu_rate, out = run(ROOT / "opal_b200" / "driver" / "opal_b200_driver", work / name / "gpu")
        except RuntimeError as err:
            line = {"deck": name, "error": str(err)[-600:]}
            print(json.dumps(line), flush=True)
            lines.append(line)
            continue
        n_out = max(int(p.name.split("_")[0]) for p in (work / name / "gpu").glob("*_energy.dat"))
        line = {"deck": name, "gpu_wall_s": gpu_wall, "gpu_particle_steps_per_s": gpu_rate, "outputs": n_out,
                "final_energies_J": energies(work / name / "gpu" / f"{n_out}_energy.dat"),
                "fits_files": len(list((work / name / "gpu").glob("*.fits")))}
        if args.ref_outputs > 0:
            env = dict(os.environ, OPAL_REF_THREADS=str(args.ref_threads))
            ref_wall, ref_rate, _ = run(ROOT / "oracle" / "opal_ref_driver", work / name / "ref", extra=("--outputs", str(args.ref_outputs)), env=env)
            k = args.ref_outputs
            eg, er = energies(work / name / "gpu" / f"{k}_energy.dat"), energies(work / name / "ref" / f"{k}_energy.dat")
            line.update({"ref_outputs": k, "ref_wall_s": ref_wall, "ref_particle_steps_per_s": ref_rate, "ref_threads": args.ref_threads,
                         "energies_at_ref_output": {"gpu": eg, "oracle": er},
                         "rel_diff": {q: (abs(eg[q] - er[q]) / abs(er[q]) if er[q] else abs(eg[q])) for q in er}})
        print(json.dumps(line), flush=True)
        lines.append(line)
    if args.qed_seeds > 0 and not args.write_only:
        import qed_deck_stats as qs
        deck = deck_text("colliding_beams")
        seeds = list(range(1, args.qed_seeds + 1))
        t0 = time.perf_counter()
        ref_dirs = qs.run_seeds(qs.REF_DRIVER, deck, work / "qed_ref", seeds, parallel=min(len(seeds), os.cpu_count() or 1))
        t1 = time.perf_counter()
        gpu_dirs = qs.run_seeds(qs.GPU_DRIVER, deck, work / "qed_gpu", seeds, parallel=1)
        t2 = time.perf_counter()
        stats = qs.compare([qs.observables(d) for d in gpu_dirs], [qs.observables(d) for d in ref_dirs])
        line = {"deck": "colliding_beams", "qed_statistics_gpu_vs_oracle": stats, "violations": qs.verdict(stats),
                "oracle_wall_s": t1 - t0, "gpu_wall_s": t2 - t1}
        print(json.dumps(line), flush=True)
        lines.append(line)
    if lines:
        pathlib.Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        pathlib.Path(args.out).write_text("\n".join(json.dumps(l) for l in lines) + "\n")


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""Runs the three BASELINE example configurations (two_stream, hole_boring, colliding_beams: SURVEY section 8 size table)
end to end through the C++ driver on the GPU, to completion, and the same deck through the oracle build of the driver up
to the first output for a full-size parity check.  Prints one JSON line per deck (also collected in --out).

The decks are regenerated here from the parameters the reference's examples/*.yaml use (the files themselves live under
/root/reference, which does not travel to the GPU box).

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

C = 2.997925e8


def deck(sections):
    out = ["---"]
    for name, fields in sections:
        out.append(f"{name}:")
        for k, v in fields:
            out.append(f" {k}: {v}")
        out.append("")
    return "\n".join(out)


def two_stream():
    return deck([
        ("control", [("dx", 500), ("nx", 1000), ("xmin", 0.0), ("start", 0.0), ("end", 0.1), ("current_deposition", "true"), ("n_outputs", 20)]),
        ("qed", [("photon_emission", "false"), ("photon_absorption", "false")]),
        ("electrons", [("npc", 100), ("ne", 20.0), ("ux", "drift_u * (1.0 + 0.001 * nrand) * (2.0 * floor(urand - 0.5) + 1.0)"), ("uy", 0.0), ("uz", 0.0),
                       ("output", "x:px")]),
        ("ions", [("npc", 0)]),
        ("constants", [("drift_u", "2.5e-24 / (me * c)")]),
    ])


def hole_boring():
    th_e, th_i = "sqrt(kT/(m*c^2)) * nrand", "sqrt(kT/(A*mp*c^2)) * nrand"
    return deck([
        ("control", [("dx", "micro / 1000"), ("nx", 20000), ("xmin", "-10*micro"), ("start", "-20.0e-6/c"), ("end", "10.0e-6/c"),
                     ("current_deposition", "true"), ("n_outputs", 30)]),
        ("qed", [("photon_emission", "false"), ("photon_absorption", "false")]),
        ("electrons", [("npc", 100), ("ne", "density * critical(omega) * step(x,xmin,xmax)"), ("ux", th_e), ("uy", th_e), ("uz", th_e),
                       ("output", "[x:px, x:p_perp, py:pz]")]),
        ("ions", [("name", "carbon"), ("npc", 100), ("Z", "Z"), ("A", "A"), ("ni", "density * critical(omega) * step(x,xmin,xmax) / Z"),
                  ("ux", th_i), ("uy", th_i), ("uz", th_i), ("output", "[x:px, x:p_perp, py:pz]")]),
        ("laser", [("Ey", "(a0*me*c*omega/e) * gauss_pulse_re(t,x,omega,sigma)"), ("Ez", "(a0*me*c*omega/e) * gauss_pulse_im(t,x,omega,sigma)")]),
        ("constants", [("density", 100.0), ("a0", 300.0), ("omega", "2*pi*c/0.8e-6"), ("sigma", "pi * 4.0 / sqrt(ln(2.0))"), ("xmin", "0.0 * micro"),
                       ("xmax", "5.0 * micro"), ("Z", 6.0), ("A", 12.0), ("kT", "500 * eV")]),
    ])


def colliding_beams():
    ey = ">\n  (a0*m*c*omega/e)\n  *sin(omega*(t-x/c))\n  *exp(-ln(2.0)*(omega*(t-x/c))^2/(2.0*pi^2*ncycles^2))"
    return deck([
        ("control", [("dx", "0.01*micro"), ("nx", 4000), ("xmin", "-15*micro"), ("start", "-24.0e-6/c"), ("end", "6.0e-6/c"),
                     ("current_deposition", "false"), ("n_outputs", 5)]),
        ("qed", [("photon_emission", "true"), ("photon_absorption", "false"), ("photon_angle_max", "50 * milli")]),
        ("electrons", [("npc", 100), ("ne", "S * a0 * critical(omega) * step(x,xmin,xmax)"), ("ux", "-1000.0 * (1.0 + 0.01 * nrand)"), ("uy", 0.0),
                       ("uz", 0.0), ("output", "[x, chi, x:chi, energy, x:energy]")]),
        ("ions", [("npc", 0)]),
        ("photons", [("npc", 0), ("output", "[x, energy, energy:(log;energy), longitude:latitude, longitude:latitude:(energy)]")]),
        ("laser", [("Ey", ey), ("Ez", 0.0)]),
        ("constants", [("S", "1.0e-6"), ("a0", 20.0), ("omega", "2*pi*c/0.8e-6"), ("ncycles", 4.0), ("xmin", "(20 - 2.5) * micro"),
                       ("xmax", "(20 + 2.5) * micro")]),
    ])


DECKS = {"two_stream": two_stream, "hole_boring": hole_boring, "colliding_beams": colliding_beams}


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
    if lines:
        pathlib.Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        pathlib.Path(args.out).write_text("\n".join(json.dumps(l) for l in lines) + "\n")


if __name__ == "__main__":
    main()

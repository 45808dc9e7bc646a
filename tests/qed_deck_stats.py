"""Deck-level statistics for QED-on runs (photon yield, energy and angle spectra within statistical error,
>= 5 seeds).  Shared by tests/test_gpu_qed_deck.py (GPU driver vs oracle driver), tests/test_examples_verbatim.py (the
oracle against itself, on the CPU, which calibrates the thresholds) and tools/run_examples.py (the full-size deck).

Both drivers are the same C++ sources (opal_b200/driver) linked over libopal_b200.so / libopal_oracle.so and run with
`--seed S --dump`: {index}_{species}.bin holds the raw per-particle arrays at every output."""
import math
import os
import pathlib
import subprocess

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
GPU_DRIVER = ROOT / "opal_b200" / "driver" / "opal_b200_driver"
REF_DRIVER = ROOT / "oracle" / "opal_ref_driver"
FIELDS = ("cell", "x", "ux", "uy", "uz", "weight", "gamma", "chi")


def load_dump(path):
    raw = np.fromfile(path, dtype=np.float64)
    n = int(np.frombuffer(raw[:1].tobytes(), dtype=np.int64)[0])
    a = raw[1:].reshape(len(FIELDS), n)
    return {k: a[i] for i, k in enumerate(FIELDS)}


def launch(driver, deck_text, workdir, seed, extra=(), env=None):
    workdir.mkdir(parents=True, exist_ok=True)
    (workdir / "in.yaml").write_text(deck_text)
    return subprocess.Popen([str(driver), str(workdir / "in.yaml"), "--seed", str(seed), "--dump", "--quiet", *extra],
                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)


def run_seeds(driver, deck_text, workdir, seeds, parallel, extra=(), threads=1, timeout=3000):
    """one run per seed; `parallel` processes at a time (the oracle runs one thread per process)"""
    env = dict(os.environ, OPAL_REF_THREADS=str(threads))
    dirs, pending = [], []
    for s in seeds:
        d = workdir / f"seed{s}"
        dirs.append(d)
        pending.append((s, launch(driver, deck_text, d, s, extra, env)))
        if len(pending) >= parallel:
            _drain(pending, timeout)
    _drain(pending, timeout)
    return dirs


def _drain(pending, timeout):
    for s, p in pending:
        out, _ = p.communicate(timeout=timeout)
        if p.returncode != 0:
            raise RuntimeError(f"seed {s}: driver failed\n{out[-2000:]}")
    pending.clear()


def last_output(d):
    return max(int(p.name.split("_")[0]) for p in d.glob("*_energy.dat"))


def observables(d, index=None):
    """photon yield, photon energies [MeV], longitude / latitude (mod.rs:440-449), electrons' mean gamma"""
    k = last_output(d) if index is None else index
    ph, el = load_dump(d / f"{k}_photon.bin"), load_dump(d / f"{k}_electron.bin")
    mag = np.sqrt(ph["ux"] ** 2 + ph["uy"] ** 2 + ph["uz"] ** 2)
    return {"n_photons": ph["x"].size, "n_electrons": el["x"].size, "photon_weight": ph["weight"].sum(),
            "energy": 0.510999 * mag, "longitude": np.arctan2(ph["uy"], -ph["ux"]), "latitude": np.arcsin(ph["uz"] / np.maximum(mag, 1e-300)),
            "gamma_mean": float(el["gamma"].mean()), "gamma_sem": float(el["gamma"].std(ddof=1) / math.sqrt(el["gamma"].size))}


def ks_2samp(a, b):
    """two-sample Kolmogorov-Smirnov statistic D and the scaled K = D sqrt(nm/(n+m)); K > 1.95 <=> p < 1e-3"""
    a, b = np.sort(a), np.sort(b)
    allv = np.concatenate([a, b])
    d = np.abs(np.searchsorted(a, allv, side="right") / a.size - np.searchsorted(b, allv, side="right") / b.size).max()
    return d, d * math.sqrt(a.size * b.size / (a.size + b.size))


def compare(obs_a, obs_b):
    """obs_*: lists of observables(), one per seed.  Returns the statistics; see `verdict` for the thresholds."""
    na, nb = sum(o["n_photons"] for o in obs_a), sum(o["n_photons"] for o in obs_b)
    out = {"seeds": [len(obs_a), len(obs_b)], "photons": [na, nb],
           # yields: both totals are sums over the same number of beam electrons
           "yield_sigma": (na - nb) / math.sqrt(max(na + nb, 1))}
    for q in ("energy", "longitude", "latitude"):
        d, k = ks_2samp(np.concatenate([o[q] for o in obs_a]), np.concatenate([o[q] for o in obs_b]))
        out[f"ks_{q}"] = {"D": d, "K": k}
    ga, gb = np.array([o["gamma_mean"] for o in obs_a]), np.array([o["gamma_mean"] for o in obs_b])
    sem = math.sqrt(sum(o["gamma_sem"] ** 2 for o in obs_a) / len(obs_a) ** 2 + sum(o["gamma_sem"] ** 2 for o in obs_b) / len(obs_b) ** 2)
    out["gamma_mean"] = [float(ga.mean()), float(gb.mean())]
    out["gamma_sigma"] = float((ga.mean() - gb.mean()) / max(sem, 1e-300))
    return out


# Thresholds.  The photons of one electron are correlated (every emission lowers gamma and with it the later rates and
# energies), so the effective sample size of a KS test is smaller than the photon count and the yield is not strictly
# Poisson; test_examples_verbatim.py measures both statistics for the oracle against itself on disjoint seed sets and
# checks that they sit well inside these bounds.
YIELD_SIGMA_MAX = 3.0
GAMMA_SIGMA_MAX = 3.0
KS_K_MAX = 1.95       # p = 1e-3 for independent samples


def verdict(stats):
    bad = []
    if abs(stats["yield_sigma"]) > YIELD_SIGMA_MAX:
        bad.append(f"photon yield differs by {stats['yield_sigma']:.2f} sigma (Poisson) {stats['photons']}")
    if abs(stats["gamma_sigma"]) > GAMMA_SIGMA_MAX:
        bad.append(f"electron mean gamma differs by {stats['gamma_sigma']:.2f} sigma {stats['gamma_mean']}")
    for q in ("energy", "longitude", "latitude"):
        if stats[f"ks_{q}"]["K"] > KS_K_MAX:
            bad.append(f"KS on photon {q}: D = {stats[f'ks_{q}']['D']:.4g}, K = {stats[f'ks_{q}']['K']:.2f} > {KS_K_MAX}")
    return bad


def shrink_deck(text, npc):
    """the reference deck with fewer macro-electrons per cell (same physics, smaller sample): only `electrons.npc` changes"""
    head, sep, tail = text.partition("electrons:\n npc: 100\n")
    assert sep, "electrons.npc: 100 not found in the deck"
    return head + f"electrons:\n npc: {npc}\n" + tail

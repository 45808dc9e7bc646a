#!/usr/bin/env python3
"""Compares the *_energy.dat files (and, where both have them, the FITS histograms' TOTAL weight) of two runs of the same deck,
e.g. the GPU driver's and the CPU oracle driver's: python tools/compare_deck_outputs.py dir_a dir_b"""
import pathlib
import re
import sys

a, b = pathlib.Path(sys.argv[1]), pathlib.Path(sys.argv[2])


def energies(p):
    return {k: float(v) for k, v in (line.split() for line in p.read_text().splitlines())}


def fits_total(p):
    m = re.search(rb"TOTAL\s*=\s*([-+0-9.eE]+)", p.read_bytes()[:2880 * 4])
    return float(m.group(1)) if m else None


worst = {}
for fa in sorted(a.glob("*_energy.dat"), key=lambda p: int(p.name.split("_")[0])):
    fb = b / fa.name
    if not fb.exists():
        continue
    ea, eb = energies(fa), energies(fb)
    rel = {k: abs(ea[k] - eb[k]) / max(abs(eb[k]), 1e-300) for k in eb if eb[k] != 0.0 or ea[k] != 0.0}
    for k, v in rel.items():
        worst[k] = max(worst.get(k, 0.0), v)
    print(fa.name.split("_")[0].rjust(3), "  ".join(f"{k} {ea[k]:.6e} vs {eb[k]:.6e} (rel {rel[k]:.1e})" for k in rel))
for fa in sorted(a.glob("*.fits")):
    fb = b / fa.name
    if fb.exists():
        ta, tb = fits_total(fa), fits_total(fb)
        if ta is not None and tb is not None:
            print(fa.name, "TOTAL", ta, "vs", tb, "rel", abs(ta - tb) / max(abs(tb), 1e-300))
print("worst relative difference per energy:", {k: f"{v:.2e}" for k, v in worst.items()})

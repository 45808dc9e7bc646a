"""The C++ driver end to end on the GPU (opal_b200/driver/opal_b200_driver: YAML -> loader -> opal_run_steps -> FITS / grid /
energy files) against the same driver sources built over the CPU oracle (oracle/opal_ref_driver) on the same input deck.
Both load identical particles (same loader, same seed), so the outputs agree to the accumulated 1e-10-per-step parity."""
import pathlib
import subprocess

import numpy as np
import pytest

from test_driver_host import parse_fits

pytestmark = pytest.mark.gpu
ROOT = pathlib.Path(__file__).resolve().parent.parent
GPU_DRIVER = ROOT / "opal_b200" / "driver" / "opal_b200_driver"
REF_DRIVER = ROOT / "oracle" / "opal_ref_driver"
DT500 = 0.95 * 500 / 2.997925e8

TWO_STREAM = """---
control:
 dx: 500
 nx: 200
 xmin: 0.0
 start: 0.0
 end: {end}
 current_deposition: true
 n_outputs: 2

qed:
 photon_emission: false
 photon_absorption: false

electrons:
 npc: 50
 ne: 20.0
 ux: drift_u * (1.0 + 0.001 * nrand) * (2.0 * floor(urand - 0.5) + 1.0)
 uy: 0.0
 uz: 0.0
 output: [x:px, px, energy]

ions:
 npc: 0

constants:
 drift_u: 2.5e-24 / (me * c)
"""

HOLE_BORING = """---
control:
 dx: micro / 100
 nx: 600
 xmin: -2*micro
 start: -3.0e-6/c
 end: {end}
 current_deposition: true
 n_outputs: 2

qed:
 photon_emission: false
 photon_absorption: false

electrons:
 npc: 20
 ne: density * critical(omega) * step(x,xmin,xmax)
 ux: sqrt(kT/(m*c^2)) * nrand
 uy: sqrt(kT/(m*c^2)) * nrand
 uz: sqrt(kT/(m*c^2)) * nrand
 output: [x:px, x:p_perp, py:pz]

ions:
 name: carbon
 npc: 20
 Z: Z
 A: A
 ni: density * critical(omega) * step(x,xmin,xmax) / Z
 ux: sqrt(kT/(A*mp*c^2)) * nrand
 uy: sqrt(kT/(A*mp*c^2)) * nrand
 uz: sqrt(kT/(A*mp*c^2)) * nrand
 output: [x:px, energy]

laser:
 Ey: (a0*me*c*omega/e) * gauss_pulse_re(t,x,omega,sigma)
 Ez: (a0*me*c*omega/e) * gauss_pulse_im(t,x,omega,sigma)

constants:
 density: 10.0
 a0: 30.0
 omega: 2*pi*c/0.8e-6
 sigma: pi * 1.0 / sqrt(ln(2.0))
 xmin: 0.0 * micro
 xmax: 2.0 * micro
 Z: 6.0
 A: 12.0
 kT: 500 * eV
"""

# a beam meeting a laser pulse head on, with photons travelling along (colliding_beams.yaml with a photon population
# and absorption switched on; emission off so that the photon population changes through absorb only)
ABSORPTION = """---
control:
 dx: 0.01*micro
 nx: 400
 xmin: -1*micro
 start: -2.0e-6/c
 end: -2.0e-6/c + 300.5 * 0.95 * 0.01*micro / c
 current_deposition: false
 n_outputs: 2

qed:
 photon_emission: false
 photon_absorption: true

electrons:
 npc: 20
 ne: dens * critical(omega) * step(x,xmin,xmax)
 ux: -1000.0 * (1.0 + 0.01 * nrand)
 uy: 0.0
 uz: 0.0
 output: [x, energy]

ions:
 npc: 0

photons:
 npc: 20
 nph: dens * critical(omega) * step(x,xmin,xmax)
 ux: -300.0 * (1.0 + urand)
 uy: 0.3 * nrand
 uz: 0.3 * nrand
 output: [x, energy]

laser:
 Ey: (a0*m*c*omega/e) * sin(omega*(t-x/c)) * exp(-ln(2.0)*(omega*(t-x/c))^2/(2.0*pi^2*ncycles^2))
 Ez: 0.0

constants:
 dens: {dens}
 a0: 100.0
 omega: 2*pi*c/0.8e-6
 ncycles: 2.0
 xmin: 1.0 * micro
 xmax: 2.5 * micro
"""


def _run(driver, deck, tmp, name, extra=()):
    d = tmp / name
    d.mkdir()
    (d / "in.yaml").write_text(deck)
    out = subprocess.run([str(driver), str(d / "in.yaml"), *extra], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    return d


def _energies(path):
    return {k: float(v) for k, v in (line.split() for line in path.read_text().splitlines())}


def _compare_dirs(a, b, indices, fits, rel):
    for idx in indices:
        ea, eb = _energies(a / f"{idx}_energy.dat"), _energies(b / f"{idx}_energy.dat")
        for k in eb:
            assert abs(ea[k] - eb[k]) <= rel * abs(eb[k]) + 1e-300, (idx, k, ea[k], eb[k])
        ga, gb = np.loadtxt(a / f"{idx}_grid.dat"), np.loadtxt(b / f"{idx}_grid.dat")
        assert ga.shape == gb.shape and np.array_equal(ga[:, 0], gb[:, 0])
        for col in range(1, 11):
            scale = np.abs(gb[:, col]).max()
            assert np.abs(ga[:, col] - gb[:, col]).max() <= rel * scale + 1e-300, (idx, col)
        for name in fits:
            ca, da = parse_fits(a / f"{idx}_{name}.fits")
            cb, db = parse_fits(b / f"{idx}_{name}.fits")
            assert da.shape == db.shape and ca["OBJECT"] == cb["OBJECT"] and ca["BUNIT"] == cb["BUNIT"]
            for key in ("TOTAL", "CRVAL1", "CDELT1"):
                assert float(ca[key]) == pytest.approx(float(cb[key]), rel=max(rel, 1e-12)), (idx, name, key)
            # a particle within round-off of a bin edge may sit in the neighbouring bin: compare the moved weight
            assert np.abs(da - db).sum() <= 5e-3 * np.abs(db).sum(), (idx, name)


def test_two_stream_deck_gpu_vs_oracle(tmp_path):
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    deck = TWO_STREAM.format(end=repr(200.5 * DT500))
    a = _run(GPU_DRIVER, deck, tmp_path, "gpu", extra=("--full",))       # the reference's whole per-particle state
    b = _run(REF_DRIVER, deck, tmp_path, "ref")
    _compare_dirs(a, b, (0, 1, 2), ("electron_x-px", "electron_px", "electron_energy"), rel=2e-6)
    # default for this deck (QED off, deposition on, no r/work/chi output): the lean (x, u, w) layout and the fused fast kernel
    c = _run(GPU_DRIVER, deck, tmp_path, "lean")
    _compare_dirs(c, b, (0, 1, 2), ("electron_x-px", "electron_px", "electron_energy"), rel=2e-6)


def test_hole_boring_deck_gpu_vs_oracle(tmp_path):
    """laser + absorbing boundaries, electrons and C6+ ions (BASELINE config 2, scaled down)"""
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    dt = 0.95 * 1.0e-8 / 2.997925e8
    deck = HOLE_BORING.format(end=repr(-3.0e-6 / 2.997925e8 + 240.5 * dt))
    a = _run(GPU_DRIVER, deck, tmp_path, "gpu", extra=("--full",))
    b = _run(REF_DRIVER, deck, tmp_path, "ref")
    fits = ("electron_x-px", "electron_x-p_perp", "electron_py-pz", "carbon_x-px", "carbon_energy")
    _compare_dirs(a, b, (0, 1), fits, rel=1e-5)
    lean = _run(GPU_DRIVER, deck, tmp_path, "lean")                      # automatic: lean electrons (fast kernel) + lean ions
    _compare_dirs(lean, b, (0, 1), fits, rel=1e-5)
    # after the pulse has hit the plasma round-off grows exponentially: conservation-level agreement
    ea, eb = _energies(a / "2_energy.dat"), _energies(b / "2_energy.dat")
    for k in ("em_field", "electrons", "ions"):
        assert abs(ea[k] - eb[k]) <= 1e-3 * abs(eb[k]), (k, ea[k], eb[k])


@pytest.mark.parametrize("name,steps,extra,fits,rel", [
    ("two_stream", 400, (), ("electron_x-px",), 2e-6),
    ("hole_boring", 100, (), ("electron_x-px", "electron_x-p_perp", "electron_py-pz", "carbon_x-px", "carbon_py-pz"), 1e-6),
    # emission on: the two sides draw different random numbers, but in the first 2 x 300 steps the beam is still 10 um away
    # from the pulse (chi = 0: no photon, tau only counts down), so the run is deterministic up to here
    ("colliding_beams", 300, (), ("electron_x", "electron_energy", "electron_x-energy"), 1e-6),
])
def test_reference_example_files_unmodified(tmp_path, name, steps, extra, fits, rel):
    """BASELINE configs 1-3: examples/*.yaml are the reference's files byte for byte (test_examples_verbatim.py); both
    drivers read THE FILE, full size (1e5 / 1e6 / 5e4 particles), shortened only by --steps (steps between outputs)."""
    import os
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    deck = (ROOT / "examples" / f"{name}.yaml").read_text()
    os.environ.setdefault("OPAL_REF_THREADS", str(os.cpu_count() or 1))
    args = ("--steps", str(steps), "--outputs", "2", *extra)
    a = _run(GPU_DRIVER, deck, tmp_path, "gpu", extra=args)
    b = _run(REF_DRIVER, deck, tmp_path, "ref", extra=args)
    _compare_dirs(a, b, (0, 1, 2), fits, rel=rel)


def test_two_ranks_two_gpus(tmp_path):
    """`--ranks 2`: one process per GPU (the reference: mpirun -n 2), x-slabs from GridDesign, NCCL ring in the library, host
    collectives for the outputs.  Each rank seeds its loader with its rank id (main.rs:53), so the plasma differs from the
    one-rank run in its random draws only: totals are exact, energies agree statistically."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    deck = TWO_STREAM.format(end=repr(200.5 * DT500))
    one = _run(GPU_DRIVER, deck, tmp_path, "one")
    two = _run(GPU_DRIVER, deck, tmp_path, "two", extra=("--ranks", "2"))
    for idx in (0, 1, 2):
        g1, g2 = np.loadtxt(one / f"{idx}_grid.dat"), np.loadtxt(two / f"{idx}_grid.dat")
        assert g1.shape == g2.shape == (200, 11) and np.array_equal(g1[:, 0], g2[:, 0])      # rank 0 gathered both slabs, in order
        c1, d1 = parse_fits(one / f"{idx}_electron_x-px.fits")
        c2, d2 = parse_fits(two / f"{idx}_electron_x-px.fits")
        assert d1.shape == d2.shape and float(c2["TOTAL"]) == pytest.approx(float(c1["TOTAL"]), rel=1e-12)   # no particle lost in migration
        e1, e2 = _energies(one / f"{idx}_energy.dat"), _energies(two / f"{idx}_energy.dat")
        assert abs(e2["electrons"] - e1["electrons"]) <= 1e-3 * e1["electrons"]
    # the field energy grows from noise: same order of magnitude
    assert 0.2 < _energies(two / "2_energy.dat")["em_field"] / _energies(one / "2_energy.dat")["em_field"] < 5.0


def test_absorption_deck_gpu_vs_oracle(tmp_path):
    """qed.photon_absorption: true (main.rs:246-248 -> interactions.rs:145).  At a physical density the pair probabilities
    are ~1e-15 per step: no event, so the run is deterministic and the two drivers agree like any other deck while the
    whole absorb path runs every step.  At an (unphysical) 1e17 times that density a few per cent of the photons are
    absorbed or copied within the run; the two random streams differ, so the surviving photon weight is compared
    statistically."""
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    a = _run(GPU_DRIVER, ABSORPTION.format(dens="1.0e-6"), tmp_path, "gpu")
    b = _run(REF_DRIVER, ABSORPTION.format(dens="1.0e-6"), tmp_path, "ref")
    for idx in (0, 1, 2):
        ea, eb = _energies(a / f"{idx}_energy.dat"), _energies(b / f"{idx}_energy.dat")
        for k in eb:
            assert abs(ea[k] - eb[k]) <= 2e-6 * abs(eb[k]) + 1e-300, (idx, k, ea[k], eb[k])
        for name in ("photon_x", "photon_energy", "electron_energy"):
            ca, da = parse_fits(a / f"{idx}_{name}.fits")
            cb, db = parse_fits(b / f"{idx}_{name}.fits")
            assert float(ca["TOTAL"]) == pytest.approx(float(cb["TOTAL"]), rel=1e-12)
            assert np.abs(da - db).sum() <= 5e-3 * np.abs(db).sum()
    quiet = float(parse_fits(b / "2_photon_x.fits")[0]["TOTAL"]) / float(parse_fits(b / "0_photon_x.fits")[0]["TOTAL"])
    c = _run(GPU_DRIVER, ABSORPTION.format(dens="1.0e11"), tmp_path, "gpu_dense")
    d = _run(REF_DRIVER, ABSORPTION.format(dens="1.0e11"), tmp_path, "ref_dense")
    left = [float(parse_fits(r / "2_photon_x.fits")[0]["TOTAL"]) / float(parse_fits(r / "0_photon_x.fits")[0]["TOTAL"]) for r in (c, d)]
    assert left[0] < quiet - 0.01 and left[1] < quiet - 0.01, (left, quiet)        # photons were absorbed on both sides
    assert abs(left[0] - left[1]) < 0.02, (left, quiet)                            # ~3000 photons, ~200 events: sigma ~ 0.005
    ec, ed = _energies(c / "2_energy.dat"), _energies(d / "2_energy.dat")
    assert abs(ec["electrons"] - ed["electrons"]) <= 2e-3 * ed["electrons"]         # the kicks are a per-mille effect here

"""Deck-level QED parity (north_star: "QED-on runs must match photon ... spectra and yields within statistical error";
SURVEY 8c: >= 5 seeds): examples/colliding_beams.yaml -- the reference's file, only `electrons.npc` lowered from 100 to 10
so that ten runs fit a test -- through opal_b200_driver (CUDA) and through the same driver sources linked over the CPU
oracle, five loader/emission seeds each, compared at the LAST output (after the beam has crossed the pulse and radiated
~4.6 photons per electron): photon yield within 3 sigma (Poisson), two-sample KS on photon energy, longitude and latitude
(mod.rs:440-449), mean electron gamma (the radiated energy) within 3 sigma.  Thresholds and their calibration:
tests/qed_deck_stats.py, tests/test_examples_verbatim.py.  Every draw is "parity unpinned" (rand_distr / xoshiro are not
in the reference tree), so this statistical agreement IS the parity statement for emit_radiation
(interactions.rs:45-107) at deck level."""
import json
import os
import pathlib
import subprocess

import pytest

import qed_deck_stats as qs

pytestmark = pytest.mark.gpu
ROOT = pathlib.Path(__file__).resolve().parent.parent
SEEDS = [1, 2, 3, 4, 5]


def test_colliding_beams_photon_statistics_gpu_vs_oracle(tmp_path):
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    deck = qs.shrink_deck((ROOT / "examples" / "colliding_beams.yaml").read_text(), 10)
    cores = os.cpu_count() or 1
    ref_dirs = qs.run_seeds(qs.REF_DRIVER, deck, tmp_path / "ref", SEEDS, parallel=min(len(SEEDS), cores))
    gpu_dirs = qs.run_seeds(qs.GPU_DRIVER, deck, tmp_path / "gpu", SEEDS, parallel=1)
    ref = [qs.observables(d) for d in ref_dirs]
    gpu = [qs.observables(d) for d in gpu_dirs]
    assert all(qs.last_output(d) == 5 for d in ref_dirs + gpu_dirs)
    assert all(o["n_electrons"] == 5000 for o in ref + gpu)                # nobody leaves the box within the run
    assert all(o["n_photons"] > 15000 for o in ref + gpu)
    stats = qs.compare(gpu, ref)
    print("colliding_beams deck-level statistics (gpu vs oracle):", json.dumps(stats))
    out = ROOT / "gpurun_out"
    if out.is_dir():
        (out / "qed_deck_stats.json").write_text(json.dumps(stats) + "\n")
    assert qs.verdict(stats) == [], stats
    # the energy the electrons lost is the energy the photons carry (both sides, every seed): recoil bookkeeping
    for d in gpu_dirs + ref_dirs:
        e0 = {k: float(v) for k, v in (line.split() for line in (d / "0_energy.dat").read_text().splitlines())}
        e5 = {k: float(v) for k, v in (line.split() for line in (d / "5_energy.dat").read_text().splitlines())}
        lost = e0["electrons"] - e5["electrons"]
        assert lost > 0.05 * e0["electrons"]
        # emission is beamed into 1/gamma = 1 mrad, far inside the 50 mrad acceptance (interactions.rs:74-95), and the laser
        # does next to no net work on the beam: the stored photons carry what the electrons lost (oracle: 1.00002)
        assert 0.98 * lost < e5["photons"] < 1.02 * lost, (d, lost, e5["photons"])

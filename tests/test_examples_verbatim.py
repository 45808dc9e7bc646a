"""The reference's three example decks (BASELINE configs 1-3) are kept byte for byte under examples/ and are what the driver,
tools/run_examples.py and the deck-level tests read -- not re-typed strings.  CPU only: the host-side reader on the files,
the oracle build of the driver on the unmodified two_stream.yaml, and the calibration of the deck-level QED statistics
(the oracle against itself on disjoint seed sets) that tests/test_gpu_qed_deck.py applies to the CUDA path."""
import hashlib
import math
import pathlib
import subprocess

import numpy as np
import pytest

import qed_deck_stats as qs
from test_driver_host import host, parse_fits, read  # noqa: F401  (host is a fixture)

ROOT = pathlib.Path(__file__).resolve().parent.parent
EX = ROOT / "examples"
# sha256 of /root/reference/examples/*.yaml (the reference does not travel to the GPU box; the digests do)
SHA256 = {
    "two_stream.yaml": "cfaa4754c66afccef6a9496dc91326ae34a5743be03ecd64492fb3876870ce4e",
    "hole_boring.yaml": "bdfc35f220f31064000f36afb0e991dd7acff360bf87b19191874ebf1168b4b9",
    "colliding_beams.yaml": "35dc576268a76c3db1cc678fd0fb3a535ac2daf0da20b9c9e195e77f6a0c36c0",
}


@pytest.mark.parametrize("name", sorted(SHA256))
def test_example_is_the_reference_file_byte_for_byte(name):
    data = (EX / name).read_bytes()
    assert hashlib.sha256(data).hexdigest() == SHA256[name]
    ref = pathlib.Path("/root/reference/examples") / name
    if ref.exists():
        assert ref.read_bytes() == data


def test_reader_on_the_verbatim_files(host):  # noqa: F811
    c, e, me, eps0 = 2.997925e8, 1.602177e-19, 9.109383e-31, 8.854188e-12
    two = (EX / "two_stream.yaml").read_text()
    assert read(host, two, "control", "nx", kind=1)[1] == 1000 and read(host, two, "control", "n_outputs", kind=1)[1] == 20
    assert read(host, two, "control", "end")[1] == 0.1 and read(host, two, "control", "current_deposition", kind=2)[1] == 1.0
    assert read(host, two, "ions", "npc", kind=1)[1] == 0            # `ions:` is followed by a line holding one blank
    assert read(host, two, "laser", "", kind=7)[1] == 0.0            # no laser section: periodic (main.rs:88-104)
    d = 2.5e-24 / (me * c)
    assert read(host, two, "electrons", "ux", kind=5, x=0.0, y=0.75, z=2.0)[1] == d * (1.0 + 0.001 * 2.0)
    hb = (EX / "hole_boring.yaml").read_text()
    omega = 2 * math.pi * c / 0.8e-6
    assert read(host, hb, "control", "nx", kind=1)[1] == 20000 and read(host, hb, "control", "n_outputs", kind=1)[1] == 30
    assert read(host, hb, "control", "dx")[1] == 1.0e-6 / 1000
    assert read(host, hb, "ions", "name", kind=3)[2] == "carbon" and read(host, hb, "ions", "A")[1] == 12.0
    crit = eps0 * me * omega ** 2 / e ** 2
    assert read(host, hb, "ions", "ni", kind=4, x=2.5e-6)[1] == pytest.approx(100.0 * crit / 6.0, rel=1e-15)
    assert read(host, hb, "laser", "", kind=7)[1] == 1.0
    cb = (EX / "colliding_beams.yaml").read_text()
    assert read(host, cb, "control", "current_deposition", kind=2)[1] == 0.0 and read(host, cb, "qed", "photon_emission", kind=2)[1] == 1.0
    assert read(host, cb, "qed", "photon_angle_max")[1] == 50 * 1.0e-3
    assert read(host, cb, "electrons", "ne", kind=4, x=20.0e-6)[1] == pytest.approx(1.0e-6 * 20.0 * crit, rel=1e-15)
    assert read(host, cb, "electrons", "ne", kind=4, x=17.4e-6)[1] == 0.0
    assert read(host, cb, "photons", "output", kind=3)[2].split("\n") == ["x", "energy", "energy:(log;energy)", "longitude:latitude",
                                                                           "longitude:latitude:(energy)"]
    # the folded `>` scalar of colliding_beams.yaml:32-35, read from the file itself
    t, x = 1.0e-15, 2.0e-7
    phi = omega * (t - x / c)
    want = (20.0 * me * c * omega / e) * math.sin(phi) * math.exp(-math.log(2.0) * phi ** 2 / (2.0 * math.pi ** 2 * 16.0))
    assert read(host, cb, "laser", "Ey", kind=6, x=t, y=x)[1] == pytest.approx(want, rel=1e-13)
    assert read(host, cb, "laser", "Ez", kind=6, x=t, y=x)[1] == 0.0


def test_oracle_driver_runs_the_unmodified_two_stream_deck(tmp_path):
    """examples/two_stream.yaml as it stands; 150 of the 3155 steps between outputs (--steps), two outputs"""
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    (tmp_path / "in.yaml").write_bytes((EX / "two_stream.yaml").read_bytes())
    out = subprocess.run([str(qs.REF_DRIVER), str(tmp_path / "in.yaml"), "--outputs", "1", "--steps", "150"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-1500:] + out.stderr[-1500:]
    assert "Output    1 at t =" in out.stdout
    cards, data = parse_fits(tmp_path / "1_electron_x-px.fits")
    assert float(cards["TOTAL"]) == pytest.approx(20.0 * 500.0 * 1000, rel=1e-12)         # ne dx nx: nobody lost on the periodic ring
    grid = np.loadtxt(tmp_path / "1_grid.dat")
    assert grid.shape == (1000, 11)
    e0 = dict(line.split() for line in (tmp_path / "0_energy.dat").read_text().splitlines())
    e1 = dict(line.split() for line in (tmp_path / "1_energy.dat").read_text().splitlines())
    assert float(e0["em_field"]) == 0.0 and float(e1["em_field"]) > 0.0                     # the instability grows from noise
    assert float(e1["electrons"]) == pytest.approx(float(e0["electrons"]), rel=1e-3)


def test_deck_level_qed_statistics_are_calibrated():
    """colliding_beams.yaml with 10 instead of 100 macro-electrons per cell, the oracle on seeds {1,2,3} against the oracle
    on seeds {4,5,6}: the statistics the GPU test applies must sit well inside their thresholds when both sides ARE the
    same code.  (Full-size, 5 + 5 seeds, measured once: yield 0.19 sigma, KS K = 0.76 / 0.57 / 0.57, gamma 1.2 sigma.)"""
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    deck = qs.shrink_deck((EX / "colliding_beams.yaml").read_text(), 10)
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        dirs = qs.run_seeds(qs.REF_DRIVER, deck, pathlib.Path(tmp), [1, 2, 3, 4, 5, 6], parallel=6)
        obs = [qs.observables(d) for d in dirs]
    assert all(o["n_electrons"] == 5000 for o in obs) and all(o["n_photons"] > 15000 for o in obs)
    stats = qs.compare(obs[:3], obs[3:])
    assert qs.verdict(stats) == [], stats
    # sensitivity: a 10 % shift of the photon energies is caught by the same test
    shifted = [dict(o, energy=1.1 * o["energy"]) for o in obs[3:]]
    assert any("energy" in m for m in qs.verdict(qs.compare(obs[:3], shifted)))

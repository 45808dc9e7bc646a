"""Host side of the driver (opal_b200/driver): the input reader (setup.rs), the meval-compatible expression evaluator,
Rust-style number formatting, output-spec parsing (particle/mod.rs:451-540), the FITS writer (hgram.rs:394-426), and a
small end-to-end run of the driver built against the CPU oracle (oracle/opal_ref_driver).  No GPU."""
import ctypes as C
import math
import pathlib
import subprocess

import numpy as np
import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
DRV = ROOT / "opal_b200" / "driver"


@pytest.fixture(scope="module")
def host():
    subprocess.check_call(["make", "-C", str(DRV), "libopal_driver_host.so"], stdout=subprocess.DEVNULL)
    lib = C.CDLL(str(DRV / "libopal_driver_host.so"))
    lib.opal_drv_config_read.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_double, C.c_double, C.c_double,
                                         C.POINTER(C.c_double), C.c_char_p, C.c_int]
    lib.opal_drv_expr_eval.argtypes = [C.c_char_p, C.POINTER(C.c_double)]
    lib.opal_drv_format.argtypes = [C.c_int, C.c_double, C.c_char_p, C.c_int]
    lib.opal_drv_output_spec.argtypes = [C.c_char_p, C.c_char_p, C.c_long, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    lib.opal_drv_write_fits.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double]
    lib.opal_drv_rng_stream.argtypes = [C.c_ulonglong, C.c_long, C.c_void_p]
    return lib


def read(host, yaml, section, field, kind=0, arg=b"x", x=0.0, y=0.0, z=0.0):
    out, buf = C.c_double(0.0), C.create_string_buffer(4096)
    rc = host.opal_drv_config_read(yaml.encode(), section.encode(), field.encode(), kind, arg, x, y, z, C.byref(out), buf, 4096)
    return rc, out.value, buf.value.decode()


def ev(host, text):
    out = C.c_double(0.0)
    rc = host.opal_drv_expr_eval(text.encode(), C.byref(out))
    return None if rc else out.value


# the reference's own parser test, setup.rs:446-484 (text verbatim, including its indentation)
CONFIG_PARSER = """---
        control:
          dx: 0.001
          nx: 4000
          ne: sin(a * x)
          ib: a * b^3

        extra:
          dx: 160

        constants:
          a: 2.0 * pi
          b: 17.0
        """


def test_reference_config_parser_golden(host):
    assert read(host, CONFIG_PARSER, "control", "dx") == (0, 0.001, "")
    assert read(host, CONFIG_PARSER, "control", "nx", kind=1)[:2] == (0, 4000.0)
    assert read(host, CONFIG_PARSER, "control", "ib")[1] == 2.0 * math.pi * 17.0 ** 3          # assert_eq! in the reference
    assert read(host, CONFIG_PARSER, "extra", "dx")[:2] == (0, 160.0)                            # integer -> f64
    assert read(host, CONFIG_PARSER, "control", "ne", kind=4, x=0.6)[1] == math.sin(2.0 * math.pi * 0.6)


def test_config_errors_and_type_dispatch(host):
    y = "control:\n nx: 10\n dx: 1.5\n flag: true\n name: abc\n expr: 2*micro\n\nconstants:\n k: 3\n r: 2.5\n s: r + 1\n"
    assert read(host, y, "nope", "nx")[0] == 2 and 'Could not find section "nope"' in read(host, y, "nope", "nx")[2]
    assert read(host, y, "control", "zz")[0] == 3
    assert read(host, y, "control", "dx", kind=1)[0] == 4            # Real is not an integer (setup.rs:302-310)
    assert read(host, y, "control", "nx", kind=2)[0] == 4
    assert read(host, y, "control", "flag", kind=2)[:2] == (0, 1.0)
    assert read(host, y, "control", "name")[0] == 4                  # unknown variable
    assert read(host, y, "control", "expr")[1] == 2 * 1.0e-6
    # constants: integers are ignored (setup.rs:164-175), a constant cannot use another constant (setup.rs:160)
    y2 = y.replace("expr: 2*micro", "expr: k")
    assert read(host, y2, "control", "expr")[0] == 4
    assert read(host, y.replace("expr: 2*micro", "expr: r"), "control", "expr")[1] == 2.5
    assert read(host, y.replace("expr: 2*micro", "expr: s"), "control", "expr")[0] == 4
    assert read(host, y, "laser", "", kind=7)[1] == 0.0 and read(host, y, "control", "", kind=7)[1] == 1.0


def test_expression_semantics(host):
    assert ev(host, "1 + 2 * 3") == 7.0
    assert ev(host, "-2^2") == -4.0                    # unary minus binds weaker than ^
    assert ev(host, "2^-1") == 0.5
    assert ev(host, "2^3^2") == 512.0                  # right associative
    assert ev(host, "7 % 4") == 3.0
    assert ev(host, "10 - 4 - 3") == 3.0 and ev(host, "16 / 4 / 2") == 2.0
    assert ev(host, "max(1, 5, 3) + min(4, 2)") == 7.0
    assert ev(host, "sqrt(16) + abs(-3) + floor(2.7) + ceil(2.1) + round(2.5) + signum(-4)") == 4 + 3 + 2 + 3 + 3 - 1
    assert ev(host, "exp(ln(2.0))") == pytest.approx(2.0, rel=1e-15)
    assert ev(host, "atan2(1, 1)") == math.atan2(1, 1)
    assert ev(host, "3!") == 6.0
    assert ev(host, "2.5e-3 * 4") == 0.01
    assert ev(host, "pi") == math.pi and ev(host, "e") == math.e      # meval's own e; opal's context overrides it (setup.rs:140)
    for bad in ("1 +", "(1", "1)", "foo(1)", "1 2", "*3", "unknown"):
        assert ev(host, bad) is None, bad


def test_example_inputs_read_like_the_reference(host):
    """the three BASELINE example files (embedded here: /root/reference does not travel)"""
    c, e, me, eps0 = 2.997925e8, 1.602177e-19, 9.109383e-31, 8.854188e-12
    two = "---\ncontrol:\n dx: 500\n nx: 1000\n xmin: 0.0\n\nelectrons:\n npc: 100\n ne: 20.0\n ux: drift_u * (1.0 + 0.001 * nrand) * (2.0 * floor(urand - 0.5) + 1.0)\n output: x:px\n\nions:\n npc: 0\n \nconstants:\n drift_u: 2.5e-24 / (me * c)\n"
    assert read(host, two, "control", "dx")[1] == 500.0
    assert read(host, two, "electrons", "ne", kind=4, x=3.0)[1] == 20.0
    d = 2.5e-24 / (me * c)
    assert read(host, two, "electrons", "ux", kind=5, x=0.0, y=0.75, z=2.0)[1] == d * (1.0 + 0.001 * 2.0) * (2.0 * math.floor(0.75 - 0.5) + 1.0)
    assert read(host, two, "electrons", "ux", kind=5, x=0.0, y=0.25, z=2.0)[1] == d * (1.0 + 0.001 * 2.0) * (2.0 * math.floor(0.25 - 0.5) + 1.0)
    assert read(host, two, "electrons", "output", kind=3)[2] == "x:px"
    hb = ("control:\n dx: micro / 1000\n xmin: -10*micro\n start: -20.0e-6/c\n\nelectrons:\n ne: density * critical(omega) * step(x,xmin,xmax)\n"
          " ux: sqrt(kT/(m*c^2)) * nrand\n output: [x:px, x:p_perp, py:pz]\n\nions:\n name: carbon\n Z: Z\n\nlaser:\n"
          " Ey: (a0*me*c*omega/e) * gauss_pulse_re(t,x,omega,sigma)\n\nconstants:\n density: 100.0\n a0: 300.0\n omega: 2*pi*c/0.8e-6\n"
          " sigma: pi * 4.0 / sqrt(ln(2.0))\n xmin: 0.0 * micro\n xmax: 5.0 * micro\n Z: 6.0\n kT: 500 * eV\n")
    omega = 2 * math.pi * c / 0.8e-6
    assert read(host, hb, "control", "dx")[1] == 1.0e-6 / 1000
    assert read(host, hb, "control", "xmin")[1] == -10 * 1.0e-6
    assert read(host, hb, "control", "start")[1] == -20.0e-6 / c
    crit = eps0 * me * omega ** 2 / e ** 2
    assert read(host, hb, "electrons", "ne", kind=4, x=1.0e-6)[1] == pytest.approx(100.0 * crit, rel=1e-15)
    assert read(host, hb, "electrons", "ne", kind=4, x=5.0e-6)[1] == 0.0 and read(host, hb, "electrons", "ne", kind=4, x=-1e-9)[1] == 0.0
    assert read(host, hb, "electrons", "ux", kind=5, x=0.0, y=0.0, z=1.5)[1] == pytest.approx(math.sqrt(500 * e / (me * c * c)) * 1.5, rel=1e-15)
    assert read(host, hb, "electrons", "output", kind=3)[2].split("\n") == ["x:px", "x:p_perp", "py:pz"]
    assert read(host, hb, "ions", "Z")[1] == 6.0
    sigma = math.pi * 4.0 / math.sqrt(math.log(2.0))
    t, x = 1.0e-15, 2.0e-7
    phi = omega * (t - x / c)
    want = (300.0 * me * c * omega / e) * (math.sin(phi) + phi * math.cos(phi) / sigma ** 2) * math.exp(-phi ** 2 / (2 * sigma ** 2))
    assert read(host, hb, "laser", "Ey", kind=6, x=t, y=x)[1] == pytest.approx(want, rel=1e-13)
    cb = ("qed:\n photon_angle_max: 50 * milli\n\nphotons:\n output: [x, energy, energy:(log;energy), longitude:latitude, longitude:latitude:(energy)]\n\n"
          "laser:\n Ey: >\n  (a0*m*c*omega/e)\n  *sin(omega*(t-x/c))\n  *exp(-ln(2.0)*(omega*(t-x/c))^2/(2.0*pi^2*ncycles^2))\n Ez: 0.0\n\n"
          "constants:\n a0: 20.0\n omega: 2*pi*c/0.8e-6\n ncycles: 4.0\n xmin: (20 - 2.5) * micro\n")
    assert read(host, cb, "qed", "photon_angle_max")[1] == 50 * 1.0e-3
    assert read(host, cb, "photons", "output", kind=3)[2].split("\n") == ["x", "energy", "energy:(log;energy)", "longitude:latitude",
                                                                           "longitude:latitude:(energy)"]
    want = (20.0 * me * c * omega / e) * math.sin(phi) * math.exp(-math.log(2.0) * phi ** 2 / (2.0 * math.pi ** 2 * 16.0))
    assert read(host, cb, "laser", "Ey", kind=6, x=t, y=x)[1] == pytest.approx(want, rel=1e-13)
    assert read(host, cb, "laser", "Ez", kind=6, x=t, y=x)[1] == 0.0


def test_rust_number_formats(host):
    def f(kind, v):
        buf = C.create_string_buffer(400)
        host.opal_drv_format(kind, v, buf, 400)
        return buf.value.decode()
    # {:e}: shortest round-trip digits, no exponent padding
    assert [f(0, v) for v in (0.0, 1.0, 500.0, 1.5e-7, -2.5, 1.0e21, 0.1, 123456.789)] == \
        ["0e0", "1e0", "5e2", "1.5e-7", "-2.5e0", "1e21", "1e-1", "1.23456789e5"]
    assert f(0, 3.393773386715751e-5) == "3.393773386715751e-5"
    # {:.6e}
    assert [f(1, v) for v in (0.0, 1.2001353e-6, 6.350928e-15, 12345678.9)] == ["0.000000e0", "1.200135e-6", "6.350928e-15", "1.234568e7"]
    # {}: positional
    assert [f(2, v) for v in (0.0, 1.0, 0.00015, -12.5, 1.0e21, 1.0e-7)] == ["0", "1", "0.00015", "-12.5", "1000000000000000000000", "0.0000001"]
    assert f(0, float("nan")) == "NaN" and f(2, float("inf")) == "inf"


def test_output_spec_parsing_and_file_names(host):
    def spec(entry, species="electron", index=3):
        info, file = C.create_string_buffer(256), C.create_string_buffer(512)
        ok = host.opal_drv_output_spec(entry.encode(), b"out", index, species.encode(), info, 256, file, 512)
        return ok, info.value.decode().split(), file.value.decode()
    ok, info, file = spec("x:px")
    assert ok and info[:4] == ["2", "0", "3", "0"] and info[6] == "0" and file == "!out/3_electron_x-px.fits"
    ok, info, file = spec("energy:(log;energy)", "photon")
    assert ok and info[:4] == ["1", "2", "-1", "1"] and info[6] == "1" and file == "!out/3_photon_energy_energy_log.fits"
    ok, info, file = spec("longitude:latitude:(energy)", "photon")
    assert ok and info[:4] == ["2", "9", "10", "0"] and info[6] == "1" and file == "!out/3_photon_longitude-latitude_energy.fits"
    ok, info, file = spec("x:(50;auto)")
    assert ok and info[3] == "2" and info[4] == "50" and file == "!out/3_electron_x_auto.fits"          # FixedNumber; weight name != "weight" is appended
    ok, info, file = spec("px:(0.5;weight)")
    assert ok and info[3] == "3" and float(info[5]) == 0.5 and file == "!out/3_electron_px.fits"
    assert spec("chi")[0] == 1 and spec("foo")[0] == 0 and spec("x:px:py")[0] == 0 and spec("x:(energy2)")[0] == 0


def parse_fits(path):
    b = pathlib.Path(path).read_bytes()
    assert len(b) % 2880 == 0
    cards, n = {}, 0
    while True:
        block = b[n:n + 2880].decode("ascii")
        n += 2880
        done = False
        for i in range(0, 2880, 80):
            c = block[i:i + 80]
            if c.startswith("END"):
                done = True
                break
            if c[8:10] == "= ":
                v = c[10:].strip()
                cards[c[:8].strip()] = v[1:v.index("'", 1)].rstrip() if v.startswith("'") else v.split("/")[0].strip()
        if done:
            break
    shape = [int(cards[f"NAXIS{k + 1}"]) for k in range(int(cards["NAXIS"]))]
    data = np.frombuffer(b[n:n + 8 * int(np.prod(shape))], dtype=">f8").reshape(shape[::-1])
    return cards, data


def test_fits_writer_roundtrip(host, tmp_path):
    bins = np.array([4, 3], dtype=np.int64)
    cts = np.arange(12, dtype=np.float64) * 1.5 - 2.0
    mn, sz = np.array([-1.0, 10.0]), np.array([0.5, 2.0])
    p = tmp_path / "h.fits"
    p.write_bytes(b"stale")
    assert host.opal_drv_write_fits(("!" + str(p)).encode(), 2, bins.ctypes.data, mn.ctypes.data, sz.ctypes.data, cts.ctypes.data, 42.5) == 0
    cards, data = parse_fits(p)
    assert cards["SIMPLE"] == "T" and cards["BITPIX"] == "-64" and cards["NAXIS"] == "2"
    assert (cards["NAXIS1"], cards["NAXIS2"]) == ("3", "4")          # fitsio reverses the row-major dimensions (hgram.rs:396-399)
    assert float(cards["CRPIX1"]) == 1.0 and float(cards["CRVAL1"]) == -0.75 and float(cards["CDELT1"]) == 0.5
    assert float(cards["CRVAL2"]) == 11.0 and float(cards["CDELT2"]) == 2.0
    assert cards["CNAME1"] == "x" and cards["CUNIT2"] == "MeV/c" and cards["BUNIT"] == "1/(m.MeV/c)" and cards["OBJECT"] == "hgram/density/x_px"
    assert float(cards["TOTAL"]) == 42.5 and float(cards["DATAMIN"]) == -2.0 and float(cards["DATAMAX"]) == 14.5
    assert np.array_equal(data.ravel(), cts)


SMALL = """---
control:
 dx: 500
 nx: 64
 xmin: 0.0
 start: 0.0
 end: {end}
 current_deposition: true
 n_outputs: 2

qed:
 photon_emission: false
 photon_absorption: false

electrons:
 npc: 20
 ne: 20.0 * step(x, 0, 16000)
 ux: drift_u * (1.0 + 0.001 * nrand) * (2.0 * floor(urand - 0.5) + 1.0)
 uy: 0.0
 uz: 0.01 * nrand
 output: [x:px, energy, "px:(8;energy)", bogus]

ions:
 npc: 0

constants:
 drift_u: 2.5e-24 / (me * c)
"""


def test_oracle_driver_end_to_end(tmp_path):
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    dt = 0.95 * 500 / 2.997925e8
    (tmp_path / "in.yaml").write_text(SMALL.format(end=repr(20.5 * dt)))
    out = subprocess.run([str(ROOT / "oracle" / "opal_ref_driver"), str(tmp_path / "in.yaml")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    # 20 steps, 2 outputs -> 10 steps between outputs; files 0, 1, 2 (main.rs:186-187, 216-276)
    for idx in (0, 1, 2):
        for name in ("grid.dat", "energy.dat", "electron_x-px.fits", "electron_energy.fits", "electron_px_energy.fits"):
            assert (tmp_path / f"{idx}_{name}").exists(), (idx, name)
    assert not list(tmp_path.glob("*bogus*")) and not (tmp_path / "3_energy.dat").exists()
    rows = np.loadtxt(tmp_path / "2_grid.dat")
    assert rows.shape == (64, 11) and np.array_equal(rows[:, 0], 500.0 * np.arange(64))
    en = dict(line.split() for line in (tmp_path / "2_energy.dat").read_text().splitlines())
    assert set(en) == {"em_field", "electrons", "ions", "photons"} and float(en["electrons"]) > 0 and en["ions"] == "0.000000e0"
    cards, data = parse_fits(tmp_path / "0_electron_x-px.fits")
    n = 32 * 20                                                         # plasma in the cells with 0 <= x < 16000
    nb = math.ceil(2 * n ** (1 / 3))
    assert data.shape == (nb, nb) and float(cards["TOTAL"]) == pytest.approx(20.0 * 500 * 32, rel=1e-12)
    # density x bin volume sums to the total weight minus the particles holding a maximum (dropped, hgram.rs:253)
    vol = float(cards["CDELT1"]) * float(cards["CDELT2"])
    assert data.sum() * vol == pytest.approx(float(cards["TOTAL"]), rel=2 * 2 / n + 1e-9)
    cards, data = parse_fits(tmp_path / "1_electron_px_energy.fits")
    assert data.shape == (8,) and cards["OBJECT"] == "hgram/density/px"
    # a missing input file is the reference's MissingFile error
    bad = subprocess.run([str(ROOT / "oracle" / "opal_ref_driver"), str(tmp_path / "none.yaml")], capture_output=True, text=True)
    assert bad.returncode == 1 and "Unable to open configuration file." in bad.stderr


def test_rng_stream_matches_the_oracles_restatement(host, oracle):
    """two independent statements of xoshiro256** / SplitMix64 seeding (driver loader and oracle) agree"""
    a, b = np.zeros(64), np.zeros(64)
    host.opal_drv_rng_stream(7, 64, a.ctypes.data)
    oracle.cdll.opal_ref_rng_stream.argtypes = [C.c_uint64, C.c_int64, C.c_void_p]
    oracle.cdll.opal_ref_rng_stream(7, 64, b.ctypes.data)
    assert np.array_equal(a, b) and 0.0 <= a.min() and a.max() < 1.0


def test_yaml_subset_features(host):
    y = ("# a comment line\r\n---\r\ncontrol:   # trailing comment\r\n  dx: 2.5   # metres\r\n  label: \"x:px # not a comment\"\r\n"
         "  outs: [x,\r\n         px:py,   # inside a flow sequence\r\n         \"energy:(log;energy)\"]\r\n"
         "  lit: |\r\n    1 +\r\n    2\r\n  fold: >-\r\n    3 *\r\n    4\r\n  hex: 0x10\r\n  neg: -7\r\n  sci: 1e3\r\n  yes: true\r\n...\r\n")
    assert read(host, y, "control", "dx")[1] == 2.5
    assert read(host, y, "control", "label", kind=3)[2] == "x:px # not a comment"
    assert read(host, y, "control", "outs", kind=3)[2].split("\n") == ["x", "px:py", "energy:(log;energy)"]
    assert read(host, y, "control", "lit")[1] == 3.0                 # "1 +\n2\n" evaluates as an expression
    assert read(host, y, "control", "fold")[1] == 12.0
    assert read(host, y, "control", "hex", kind=1)[1] == 16.0 and read(host, y, "control", "neg")[1] == -7.0
    assert read(host, y, "control", "neg", kind=1)[0] == 4           # usize::try_from(-7) fails (setup.rs:312-318)
    assert read(host, y, "control", "sci")[1] == 1000.0 and read(host, y, "control", "sci", kind=1)[0] == 4   # "1e3" is a Real, not an Integer
    assert read(host, y, "control", "yes", kind=2)[1] == 1.0
    assert read(host, "not yaml at all", "control", "dx")[0] == 1    # MissingFile, as setup.rs:96-99 maps load errors


def test_cpp_grid_design_equals_the_python_mirror(host):
    from opal_b200.slabs import GridDesign
    host.opal_drv_grid_design.argtypes = [C.c_int, C.c_int, C.c_long, C.c_double, C.c_double, C.c_long, C.c_char_p, C.c_void_p, C.c_void_p]
    cases = [(False, 4, 1003, "1"), (False, 8, 20, "1"), (True, 3, 600, "1 + 0*x"), (True, 4, 2000, "step(x, 0, 2.5e-6) + 0.01"),
             (True, 2, 64, "exp(-x*x*1.0e10)"), (True, 1, 40, "1")]
    for balanced, np_, size, ne in cases:
        nx, off = (C.c_long * np_)(), (C.c_double * np_)()
        expr = ne.replace("step(x, 0, 2.5e-6)", "max(0, min(1, floor(1 + x*1.0e9) * floor(1 + (2.5e-6 - x)*1.0e9)))")   # meval has no step(); opal adds it in Config
        xmin, dx = -1.0e-6, 5.0e-9
        assert host.opal_drv_grid_design(int(balanced), np_, size, xmin, dx, 8, expr.encode(), nx, off) == 0
        import math
        env = {"x": 0.0, "exp": math.exp, "floor": math.floor, "max": max, "min": min}
        f = lambda x: eval(expr, {"__builtins__": {}}, dict(env, x=x))
        want = GridDesign.balanced(np_, size, xmin, dx, f) if balanced else GridDesign.unbalanced(np_, size, xmin, dx)
        assert list(nx) == list(want.nx), (balanced, np_, size, ne)
        assert list(off) == [float(v) for v in want.offset]


def test_loader_draw_order_and_moments(host):
    """Population::new (mod.rs:172-203): npc particles in every cell whose density is positive, weight = ne dx / npc,
    u from the deck's expressions of (x, urand, nrand); statistics of the draws."""
    host.opal_drv_load.restype = C.c_long
    host.opal_drv_load.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_long, C.c_long, C.c_double, C.c_double, C.c_ulonglong, C.c_int, C.c_long] + [C.c_void_p] * 7
    y = ("electrons:\n ne: 3.0 * step(x, 1000, 4000)\n ux: 0.5 * nrand\n uy: urand\n uz: x / 1000\n")
    cap = 200000
    cell = np.zeros(cap, dtype=np.int64)
    x, ux, uy, uz, w, tau = (np.zeros(cap) for _ in range(6))
    n = host.opal_drv_load(y.encode(), b"electrons", b"ne", 400, 100, 0.0, 100.0, 5, 1, cap, *(a.ctypes.data for a in (cell, x, ux, uy, uz, w, tau)))
    # cell centres 50, 150, ...: density positive where 1000 <= centre < 4000 -> cells 10..39
    assert n == 30 * 400
    cell, x, ux, uy, uz, w, tau = (a[:n] for a in (cell, x, ux, uy, uz, w, tau))
    assert np.array_equal(np.unique(cell), np.arange(10, 40)) and np.all(np.bincount(cell)[10:40] == 400)
    assert np.all(w == 3.0 * 100.0 / 400) and np.all((x >= 0) & (x < 1))
    assert np.allclose(uz, (cell + x) * 100.0 / 1000, rtol=1e-14)                 # real_x = xmin + (c + x) dx
    assert abs(ux.mean()) < 4 * 0.5 / np.sqrt(n) and abs(ux.std() / 0.5 - 1) < 0.03
    assert abs(uy.mean() - 0.5) < 4 / np.sqrt(12 * n) and uy.min() >= 0 and uy.max() < 1
    assert abs(tau.mean() - 1.0) < 0.05 and tau.min() > 0                          # Exp1 optical depths
    # same seed -> same stream; the rank id seeds it (main.rs:53)
    keep = (np.zeros(cap, dtype=np.int64),) + tuple(np.zeros(cap) for _ in range(6))
    n2 = host.opal_drv_load(y.encode(), b"electrons", b"ne", 400, 100, 0.0, 100.0, 5, 1, cap, *(a.ctypes.data for a in keep))
    assert n2 == n and np.array_equal(keep[2][:n], ux) and np.array_equal(keep[6][:n], tau)
    n3 = host.opal_drv_load(y.encode(), b"electrons", b"ne", 400, 100, 0.0, 100.0, 6, 1, cap, *(a.ctypes.data for a in keep))
    assert n3 == n and not np.array_equal(keep[2][:n], ux)


def test_shipped_example_decks(host, tmp_path):
    """examples/*.yaml (our own decks for BASELINE configs 4 and 5, in opal's schema): the two full-size ones parse to the
    bench workload; the small one runs end to end through the driver built over the CPU oracle and conserves energy."""
    c, e, me = 2.997925e8, 1.602177e-19, 9.109383e-31
    ex = ROOT / "examples"
    big = (ex / "thermal_bench.yaml").read_text()
    assert read(host, big, "control", "nx")[1] == 1.0e6 and read(host, big, "electrons", "npc")[1] == 1000.0
    assert read(host, big, "control", "end")[1] == pytest.approx(110 * 0.95 * 500 / c, rel=1e-15)
    u_th = math.sqrt(500 * e / (me * c * c))
    assert u_th == pytest.approx(0.03128, rel=1e-3)                       # bench.py's U_TH, hole_boring's kT
    assert read(host, big, "electrons", "uy", kind=5, x=0.0, y=0.3, z=-2.0)[1] == pytest.approx(-2.0 * u_th, rel=1e-15)
    assert read(host, big, "control", "balance", kind=2)[:2] == (0, 0.0)
    ws = (ex / "two_stream_weak_scaling.yaml").read_text()
    assert read(host, ws, "control", "nx")[1] == 8.0e6
    d = 2.5e-24 / (me * c)
    assert read(host, ws, "electrons", "ux", kind=5, x=0.0, y=0.75, z=1.0)[1] == d * 1.001
    assert read(host, ws, "electrons", "ux", kind=5, x=0.0, y=0.25, z=1.0)[1] == -d * 1.001

    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    (tmp_path / "thermal_small.yaml").write_text((ex / "thermal_small.yaml").read_text())
    out = subprocess.run([str(ROOT / "oracle" / "opal_ref_driver"), str(tmp_path / "thermal_small.yaml")], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    en = []
    for idx in (0, 1, 2):                                                   # 220 steps, 2 outputs
        for name in ("grid.dat", "electron_x-px.fits", "electron_energy.fits", "electron_px-py.fits"):
            assert (tmp_path / f"{idx}_{name}").exists(), (idx, name)
        row = dict(line.split() for line in (tmp_path / f"{idx}_energy.dat").read_text().splitlines())
        en.append(float(row["em_field"]) + float(row["electrons"]))
    assert not (tmp_path / "3_energy.dat").exists()
    assert en[0] > 0 and abs(en[2] / en[0] - 1.0) < 1e-4                    # a quiet thermal plasma keeps its energy


COLD_BEAM = """constants:
  cfl: 0.95
electrons:
  npc: 40
  ne: 20.0
  ux: 0.3
  uy: 0.0
  uz: 0.0
  output: [x:px, energy]
ions:
  npc: 0
qed:
  photon_absorption: false
  photon_emission: false
control:
  nx: 120
  dx: 500.0
  xmin: 0.0
  balance: false
  current_deposition: true
  start: 0.0
  end: 60.5 * cfl * 500.0 / c
  n_outputs: 2
"""

LASER_SLAB = """constants:
  omega: 2 * pi * c / 0.8e-6
  a0: 5.0
electrons:
  npc: 16
  ne: 10.0 * critical(omega) * step(x, 1.0 * micro, 2.0 * micro)
  ux: 0.01 * nrand
  uy: 0.01 * nrand
  uz: 0.0
  output: [x:px, energy]
ions:
  npc: 8
  name: carbon
  Z: 6.0
  A: 12.0
  ni: 10.0 * critical(omega) * step(x, 1.0 * micro, 2.0 * micro) / 6.0
  ux: 0.0
  uy: 0.0
  uz: 0.0
  output: [x]
laser:
  Ey: (a0 * me * c * omega / e) * sin(omega * (t - x / c))
  Ez: 0.0
qed:
  photon_absorption: false
  photon_emission: true
  photon_angle_max: 0.5
photons:
  npc: 0
  output: [energy]
control:
  nx: 150
  dx: 0.02 * micro
  xmin: 0.0
  balance: {balance}
  current_deposition: true
  start: 0.0
  end: 120.5 * 0.95 * 0.02 * micro / c
  n_outputs: 2
"""


def _drive(tmp_path, name, deck, *extra):
    d = tmp_path / name
    d.mkdir()
    (d / "in.yaml").write_text(deck)
    out = subprocess.run([str(ROOT / "oracle" / "opal_ref_driver"), str(d / "in.yaml"), "--quiet", *extra], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    en = [dict(line.split() for line in (d / f"{i}_energy.dat").read_text().splitlines()) for i in range(3)]
    return d, [{k: float(v) for k, v in e.items()} for e in en]


def test_ranks_with_host_staged_exchange(tmp_path):
    """`--ranks N` over a library without a communicator (the CPU oracle answers OPAL_ENCCL): the driver moves leavers and halo
    strips itself through opal_migrants_take/give and opal_halo_take/give (ranks.h, the path of a host that keeps its own
    p2p).  A cold beam (every particle the same gamma, 640 particles crossing each slab edge; 120 cells, since the
    reference's even split drops a remainder, grid/mod.rs:128) makes the energies
    independent of the per-rank loader seeds (main.rs:53), so 1, 2 and 3 ranks must agree to the printed digits."""
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    d1, e1 = _drive(tmp_path, "one", COLD_BEAM)
    g1 = np.loadtxt(d1 / "2_grid.dat")
    for n in (2, 3):
        dn, en = _drive(tmp_path, f"r{n}", COLD_BEAM, "--ranks", str(n))
        for a, b in zip(e1, en):
            assert b["electrons"] == pytest.approx(a["electrons"], rel=2e-6)          # nobody lost or doubled at a slab edge
            assert b["em_field"] == pytest.approx(a["em_field"], rel=1e-4, abs=1e-30)
        gn = np.loadtxt(dn / "2_grid.dat")
        assert gn.shape == g1.shape and np.array_equal(gn[:, 0], g1[:, 0])            # slabs gathered in rank order (yee.rs:755-760)
        assert gn[:, 5].mean() == pytest.approx(g1[:, 5].mean(), rel=1e-6)            # the bulk E_x the uniform current drives
        assert np.abs(gn[:, 5] - g1[:, 5]).max() < 1e-2 * abs(g1[:, 5].mean())       # no seam at the slab edges
        c1, h1 = parse_fits(d1 / "2_electron_x-px.fits")
        cn, hn = parse_fits(dn / "2_electron_x-px.fits")
        assert float(cn["TOTAL"]) == pytest.approx(float(c1["TOTAL"]), rel=1e-12)


@pytest.mark.parametrize("balance", ["true", "false"])
def test_ranks_laser_deck_with_ions_and_emission(tmp_path, balance):
    """laser / absorbing boundaries on the outer ranks only (yee.rs:278-292), a slab target that leaves some ranks empty,
    ions, photon emission: three host-staged ranks run it and stay statistically on the one-rank run"""
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    deck = LASER_SLAB.replace("{balance}", balance)
    _, e1 = _drive(tmp_path, "one", deck)
    _, e3 = _drive(tmp_path, "three", deck, "--ranks", "3")
    assert e1[2]["em_field"] > 0 and e1[2]["electrons"] > 10 * e1[0]["electrons"]     # the laser arrived and heated the slab
    assert e3[2]["em_field"] == pytest.approx(e1[2]["em_field"], rel=0.02)            # injected pulse: deterministic up to the plasma's response
    assert e3[2]["electrons"] == pytest.approx(e1[2]["electrons"], rel=0.25)          # different loader seeds per rank: statistical
    assert e3[2]["ions"] == pytest.approx(e1[2]["ions"], rel=0.35)
    assert e3[0]["electrons"] == pytest.approx(e1[0]["electrons"], rel=0.2)


def test_reference_time_format(host):
    """setup.rs:487-492 (the reference's own test) plus the neighbouring SI prefixes and PrettyDuration (setup.rs:400-414)"""
    def f(kind, v):
        buf = C.create_string_buffer(64)
        host.opal_drv_format(kind, v, buf, 64)
        return buf.value.decode()
    assert f(3, 2.6e-4) == "  260.00 μs"
    assert f(3, 1.5) == "    1.50  s" and f(3, 250.0) == "  250.00  s"
    assert f(3, 3.0e-3) == "    3.00 ms" and f(3, 9.995e-10) == "  999.50 ps"
    assert f(3, 4.2e-14) == "   42.00 fs" and f(3, 7.0e-17) == "   70.00 as" and f(3, 1.0e-21) == "    0.00 as"
    assert f(3, -2.0e-5) == "  -20.00 μs" and f(3, 0.0) == "    0.00 as"
    assert f(4, 59.9) == "00:00:59" and f(4, 3725.0) == "01:02:05" and f(4, 93784.2) == "1d 02:03:04"


SINGLE_POINT = """electrons:
  npc: 1
  ne: 20.0 * step(x, 1000.0, 1500.0)
  ux: 0.5
  uy: 0.3
  uz: 0.0
  output: [energy:p_perp, "energy:p_perp:(log;weight)", x]
ions:
  npc: 1
  name: proton
  Z: 1.0
  A: 1.0
  ni: 0.0
  ux: 0.0
  uy: 0.0
  uz: 0.0
  output: [x:px]
qed:
  photon_absorption: false
  photon_emission: false
control:
  nx: 16
  dx: 500.0
  xmin: 0.0
  balance: false
  current_deposition: true
  start: 0.0
  end: 2.5 * 0.95 * 500.0 / c
  n_outputs: 1
"""


def test_reference_histogram_smoke_cases(tmp_path):
    """the reference's three #[ignore]d histogram tests (hgram.rs:435-510) through the driver: a single data point gives a
    histogram (Automatic and LogScaled bins) that writes as FITS; an empty population gives None, i.e. no file"""
    subprocess.check_call(["make", "-C", str(ROOT / "oracle")], stdout=subprocess.DEVNULL)
    (tmp_path / "in.yaml").write_text(SINGLE_POINT)
    out = subprocess.run([str(ROOT / "oracle" / "opal_ref_driver"), str(tmp_path / "in.yaml"), "--quiet"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    for idx in (0, 1):
        for name in ("electron_energy-p_perp.fits", "electron_energy-p_perp_log.fits", "electron_x.fits"):
            cards, data = parse_fits(tmp_path / f"{idx}_{name}")
            assert float(cards["TOTAL"]) == 20.0 * 500.0                          # one macro-particle carrying the cell's electrons
            assert data.shape == ((1, 1) if "p_perp" in name else (1,))          # min == max: one bin of unit volume (hgram.rs:131-134, 202)
            assert data.flat[0] > 0
            if "log" not in name:
                assert data.flat[0] == 20.0 * 500.0
        assert not list(tmp_path.glob(f"{idx}_proton*"))                          # empty_2d: generate_* returns None

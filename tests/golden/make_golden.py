"""Generates the committed fixtures of tests/golden/ from the CPU oracle (oracle/libopal_oracle.so).

The reference holds no golden files for this path and, being Rust, cannot be run here,
so these vectors are outputs of the oracle *after* it passed the reference's 13 known-answer tests
(tests/test_oracle_kat.py).  They freeze that pinned state: a later edit of the oracle that moves a
result is caught by tests/test_golden.py on the CPU, and the CUDA path is compared with the same
arrays on the GPU box, where /root/reference does not exist.

    python tests/golden/make_golden.py        # rewrites tests/golden/*.npz

Cases (all seeded, numpy default_rng):
  two_stream_small   periodic, 64 cells x 16 e-, 25 fused steps (main.rs:238-267), deposition on
  laser_slab         laser / absorbing boundaries, e- + C6+ slab, 60 call-by-call steps, the pulse reaches the slab
  single_push        one Vay push (electron.rs:268) and one Boris push (ion.rs:168) in random static fields
  qed_functions      rate (photon_emission.rs:59) over both asymptotic branches and the table, Airy Ai (airy.rs:19) over
                     its five ranges, absorption / stimulated-emission cross sections (photon_absorption.rs:17,
                     stimulated_emission.rs:17) for 1500 seeded photon-electron pairs; keys q_*: compared element-wise
  absorb_events      absorb (interactions.rs:145): surviving photons, remaining optical depths and kicked electrons of a seeded
                     scene (oracle regression only; the CUDA path meets the live oracle on it in tests/test_gpu_absorb.py)
"""
import ctypes
import math
import os
import pathlib
import sys

import numpy as np

HERE = pathlib.Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from opal_b200.abi import BDY_ABSORBING, BDY_LASER, ELECTRON, ION, PHOTON, OpalABI, Slab  # noqa: E402
from opal_b200.constants import ELECTRON_CHARGE, ELECTRON_MASS, SPEED_OF_LIGHT, VACUUM_PERMITTIVITY  # noqa: E402
import pic_cases as pc  # noqa: E402

C = SPEED_OF_LIGHT
PFIELDS = ("cell", "weight", "x", "prev_x", "ux", "uy", "uz", "gamma")


def _pack(prefix, d):
    d = pc.canon(d)
    return {f"{prefix}_{k}": d[k] for k in PFIELDS}


def two_stream_small(abi):
    nx, npc, dx, nsteps = 64, 16, 500.0, 25
    s = Slab(abi, nx, 0.0, dx)
    pc.upload(s, ELECTRON, pc.two_stream_particles(nx, npc, dx, seed=21, thermal=0.03))
    s.run_steps(nsteps)
    out = dict(fields=s.fields(), energy=np.array(s.energy()))
    out.update(_pack("e", s.download(ELECTRON)))
    return out


def laser_pulse():
    micro = 1.0e-6
    omega = 2.0 * math.pi * C / (0.8 * micro)
    e0 = 10.0 * ELECTRON_MASS * C * omega / abs(ELECTRON_CHARGE)
    ncrit = VACUUM_PERMITTIVITY * ELECTRON_MASS * omega ** 2 / ELECTRON_CHARGE ** 2

    def ly(t, x):
        phi = omega * (t - x / C)
        return e0 * math.sin(phi) * math.exp(-(phi / (2 * math.pi * 4.0)) ** 2)

    return micro, ncrit, ly, (lambda t, x: 0.0)


def laser_slab(abi):
    micro, ncrit, ly, lz = laser_pulse()
    nx, dx, nsteps = 160, 0.02 * micro, 60
    s = Slab(abi, nx, -1.0 * micro, dx, None, BDY_LASER, BDY_ABSORBING)
    cells = np.arange(50, 110)
    u_th = math.sqrt(500.0 / 0.510999e6)
    s.set_ion(6.0, 12.0)
    pc.upload(s, ELECTRON, pc.thermal_particles(nx, 8, dx, u_th, seed=31, ne=20.0 * ncrit, cells=cells))
    pc.upload(s, ION, pc.thermal_particles(nx, 8, dx, 0.0, seed=32, ne=20.0 * ncrit / 6.0, cells=cells))
    t = -1.0 * micro / C
    for _ in range(nsteps):
        t = pc.step_by_calls(s, t, laser=(ly, lz), species=(ELECTRON, ION))
    out = dict(fields=s.fields(), energy=np.array(s.energy()))
    out.update(_pack("e", s.download(ELECTRON)))
    out.update(_pack("i", s.download(ION)))
    return out


def single_push_fields(shape):
    return np.random.default_rng(41).standard_normal((shape[0], 6)) * np.array([1e9, 1e9, 1e9, 1.0, 1.0, 1.0])


def single_push(abi):
    nx, npc, dx = 32, 8, 1.0e-7
    s = Slab(abi, nx, 0.0, dx)
    f = s.fields()
    f[:, 5:11] = single_push_fields(f.shape)
    s.set_fields(f)
    s.set_ion(3.0, 7.0)
    pc.upload(s, ELECTRON, pc.thermal_particles(nx, npc, dx, 0.4, seed=42, ne=1.0e20))
    pc.upload(s, ION, pc.thermal_particles(nx, npc, dx, 0.01, seed=43, ne=1.0e20))
    s.advance(ELECTRON)
    s.advance(ION)
    out = {}
    out.update(_pack("e", s.download(ELECTRON)))
    out.update(_pack("i", s.download(ION)))
    return out


def qed_functions(abi):
    import kat
    chi = np.logspace(-4, 3, 400)
    x = np.concatenate([np.linspace(0.0, 2.0, 201), np.linspace(2.0, 49.99, 400)])
    ai, ok = abi.airy(x)
    assert ok.all()
    k4, p4, cg, ce = kat.pair_cases(seed=51, n=1500)
    sa, ss, oka, oks = abi.pair_cross_sections(k4, p4, cg, ce)
    return dict(q_rate=abi.rate(chi, np.full_like(chi, 1234.5)), q_airy=ai, q_sigma_abs=sa, q_sigma_st=ss,
                q_abs_defined=oka.astype(np.int64), q_st_defined=oks.astype(np.int64))


def absorb_events(abi):
    """absorb (interactions.rs:145) on tests/absorb_cases.py's scene with half of the photons close to absorption and the
    stimulated-emission depth switched off: which photons survive, what is left of their depths, the kicked electrons"""
    import absorb_cases as ac
    s = ac.build(abi)
    n = s.count(PHOTON)
    rng = np.random.default_rng(1)
    ac.set_depths(s, np.where(rng.random(n) < 0.5, rng.uniform(0.0, 0.3, n), 50.0), np.full(n, np.inf))
    s.absorb(0.0)
    el, ph, ta, ts = ac.snapshot(s)
    out = {"a_survivors": np.array([ph["cell"].size]), "a_tau_abs": ta}
    out.update({f"a_e_{k}": el[k] for k in ("cell", "ux", "uy", "uz", "gamma")})
    out.update({f"a_ph_{k}": ph[k] for k in ("cell", "x", "ux", "uy", "uz", "weight")})
    return out


CPU_ONLY = {"absorb_events"}      # on the GPU this scene needs OPAL_FLAG_STRICT_FP and its own tolerances: tests/test_gpu_absorb.py
CASES = dict(absorb_events=absorb_events, two_stream_small=two_stream_small, laser_slab=laser_slab, single_push=single_push, qed_functions=qed_functions)


def main():
    so = ROOT / "oracle" / "libopal_oracle.so"
    oracle = OpalABI(ctypes.CDLL(os.fspath(so)), "opal_ref_")
    for name, fn in CASES.items():
        out = fn(oracle)
        np.savez_compressed(HERE / f"{name}.npz", **out)
        print(name, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()

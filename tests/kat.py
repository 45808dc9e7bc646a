"""The reference's own known-answer tests for the hot path, restated against the C ABI so the
same checks run on the CPU oracle (opal_ref_*) and on the CUDA library (opal_*).

Sources: src/grid/yee.rs:903-1079, src/particle/electron.rs:401-436,
src/qed/photon_emission.rs:299-336 (tolerances as in the reference).
"""
import math

import numpy as np

from opal_b200.abi import (BDY_ABSORBING, BDY_CONDUCTING, BDY_INTERNAL, BDY_LASER, ELECTRON, Slab)
from opal_b200.constants import (ELECTRON_CHARGE, ELECTRON_MASS, SPEED_OF_LIGHT, VACUUM_PERMITTIVITY)

C = SPEED_OF_LIGHT


def laser_bc(abi):
    """yee.rs:903-953: inject a cos^2 pulse; EM energy = eps0 E0^2 3 pi c / omega (1e-3)."""
    xmin, dx = -10.0e-6, 1.0e-6 / 100.0
    t = -15.0e-6 / C
    dt = 0.95 * dx / C
    nsteps = int(abs(t) / dt)
    omega = 2.0 * math.pi * C / 1.0e-6
    emax = 1.0e10

    def laser_y(t, z):
        phi = omega * (t - z / C)
        return emax * math.sin(phi) * math.cos(phi / 16.0) ** 2 if abs(phi) < 8.0 * math.pi else 0.0

    g = Slab(abi, 2000, xmin, dx, dt, BDY_LASER, BDY_ABSORBING)
    x2 = g.laser_x()
    for _ in range(nsteps + 1):
        g.synchronize(t, laser_y(t, x2), 0.0)
        g.field_advance()
        t += dt
    E, _ = g.fields_at([1000 + 25, 1000 + 125], [0.0, 0.0])
    em = g.energy()[0]
    target = VACUUM_PERMITTIVITY * emax ** 2 * (3.0 * math.pi * C / omega)
    g.close()
    return em, target, E[:, 1] / emax


def conducting_bc(abi):
    """yee.rs:956-1011: circularly polarised pulse reflected by a conductor (1e-3)."""
    xmin, dx = -10.0e-6, 1.0e-6 / 100.0
    t = -15.0e-6 / C
    dt = 0.95 * dx / C
    nsteps = int(35.0e-6 / (C * dt))
    omega = 2.0 * math.pi * C / 1.0e-6
    emax = 1.0e10

    def laser_y(t, z):
        phi = omega * (t - z / C)
        if abs(phi) < 8.0 * math.pi:
            return -emax * (math.sin(phi) + (2.0 / 16.0) * math.cos(phi) * math.tan(phi / 16.0)) * math.cos(phi / 16.0) ** 2
        return 0.0

    def laser_z(t, z):
        phi = omega * (t - z / C)
        if abs(phi) < 8.0 * math.pi:
            return emax * (math.cos(phi) + (2.0 / 16.0) * math.sin(phi) * math.tan(phi / 16.0)) * math.cos(phi / 16.0) ** 2
        return 0.0

    g = Slab(abi, 2000, xmin, dx, dt, BDY_LASER, BDY_CONDUCTING)
    x2 = g.laser_x()
    for _ in range(nsteps + 1):
        g.synchronize(t, laser_y(t, x2), laser_z(t, x2))
        g.field_advance()
        t += dt
    em = g.energy()[0]
    target = VACUUM_PERMITTIVITY * emax ** 2 * (193.0 / 32.0) * (math.pi * C / omega)
    g.close()
    return em, target


def periodic_bc(abi):
    """yee.rs:1014-1079: antenna current on a periodic grid; energy at 2T == energy at T (1e-6)."""
    xmin, dx = -10.0e-6, 1.0e-6 / 100.0
    t = -5.0e-6 / C
    dt = 0.95 * dx / C
    nsteps = int(10.0e-6 / (C * dt))
    omega = 2.0 * math.pi * C / 1.0e-6
    jmax = omega * VACUUM_PERMITTIVITY * 1.0e5
    g = Slab(abi, 2000, xmin, dx, dt, BDY_INTERNAL, BDY_INTERNAL)

    def drive(t):
        phi = omega * t
        if abs(phi) < 4.0 * math.pi:
            return jmax * (math.sin(phi) + (2.0 / 8.0) * math.cos(phi) * math.tan(phi / 8.0)) * math.cos(phi / 8.0) ** 2
        return 0.0

    energies = []
    for _half in range(2):
        for _ in range(nsteps):
            f = g.fields()
            f[50 * 20, 4] = drive(t)      # cell[1000].j[2]  (Cell = x, rho, j[3], E[3], B[3])
            g.set_fields(f)
            g.synchronize(t, 0.0, 0.0)
            g.field_advance()
            t += dt
        energies.append(g.energy()[0])
    g.close()
    return energies


def _single_electron(abi, u0, E, B, dx, dt, nsteps, nx=64):
    g = Slab(abi, nx, 0.0, dx, dt, BDY_INTERNAL, BDY_INTERNAL)
    f = g.fields()
    f[:, 5:8] = E
    f[:, 8:11] = B
    g.set_fields(f)
    g.upload(ELECTRON, [nx // 2], [0.0], [u0[0]], [u0[1]], [u0[2]], [1.0])
    for _ in range(nsteps):
        g.advance(ELECTRON)
    p = g.download(ELECTRON)
    g.close()
    return p


def static_magnetic_field(abi):
    """electron.rs:401-418: one gyro-period in 100 steps; gamma conserved (one-sided 1e-6)."""
    b0, u0 = 1.0, 1.0
    r_c = ELECTRON_MASS * C * u0 / (abs(ELECTRON_CHARGE) * b0)
    omega_c = C * u0 / (math.sqrt(1.0 + u0 * u0) * r_c)
    dt = 0.01 * 2.0 * math.pi / omega_c
    p = _single_electron(abi, (u0, 0.0, 0.0), (0.0, 0.0, 0.0), (0.0, 0.0, b0), 1.0, dt, 100)
    return p, math.sqrt(1.0 + u0 * u0)


def static_electric_field(abi):
    """electron.rs:424-436: gamma(t) = sqrt(1 + (e Ex t / m c)^2) after 100 steps of 1 ns."""
    e0, dt = 1.0e3, 1.0e-9
    p = _single_electron(abi, (0.0, 0.0, 0.0), (e0, 0.0, 0.0), (0.0, 0.0, 0.0), 1.0, dt, 100)
    target = math.sqrt(1.0 + (ELECTRON_CHARGE * e0 * 1.0e-7 / (ELECTRON_MASS * C)) ** 2)
    return p, target


RATE_KATS = [(0.026, 2.07935e14), (3.5, 1.58485e16), (9.98, 3.45844e16), (12.4, 4.04647e16), (403.0, 4.46834e17)]


def rate_kats(abi):
    """qed/photon_emission.rs:299-336: rate(chi, gamma=1000) against five tabulated values (1e-3)."""
    chi = np.array([c for c, _ in RATE_KATS])
    got = abi.rate(chi, np.full_like(chi, 1000.0))
    want = np.array([v for _, v in RATE_KATS])
    return got, want


# qed/special_functions/airy.rs:77-112: Ai at 0, 2, 17, 20 to 1e-12 relative; Ai(200) is None
AIRY_KATS = ((0.0, 0.3550280538878172), (2.0, 0.03492413042327438), (17.0, 7.05019729838861e-22), (20.0, 1.69167286867e-27))


def airy_kats(abi):
    x = np.array([v for v, _ in AIRY_KATS] + [200.0, -1.0, 50.0, float("nan")])
    ai, ok = abi.airy(x)
    rel = [abs(ai[i] - t) / t for i, (_, t) in enumerate(AIRY_KATS)]
    return rel, ok


def pair_cases(seed=7, n=4000):
    """photon / electron pairs spanning the branches of the two cross sections: both chi orderings, k0 above and below
    gamma, zero chi, and zbar in each of the Airy function's five ranges"""
    rng = np.random.default_rng(seed)
    gamma = 10.0 ** rng.uniform(0.5, 4.0, n)
    u = np.sqrt(gamma ** 2 - 1.0)
    ct = rng.uniform(-1.0, 1.0, n)
    ph = rng.uniform(0.0, 2.0 * np.pi, n)
    st = np.sqrt(1.0 - ct ** 2)
    p4 = np.stack([gamma, u * ct, u * st * np.cos(ph), u * st * np.sin(ph)], axis=1)
    k0 = gamma * 10.0 ** rng.uniform(-6.0, 0.3, n)
    # nearly collinear with the electron (k.p small, zbar in the series range) or at a wide angle (zbar large)
    dth = 10.0 ** rng.uniform(-5.0, 0.5, n)
    ct2 = np.cos(np.arccos(ct) + dth)
    st2 = np.sqrt(np.maximum(0.0, 1.0 - ct2 ** 2))
    k4 = np.stack([k0, k0 * ct2, k0 * st2 * np.cos(ph), k0 * st2 * np.sin(ph)], axis=1)
    chi_e = 10.0 ** rng.uniform(-3.0, 1.0, n)
    chi_g = chi_e * 10.0 ** rng.uniform(-4.0, 0.5, n)
    chi_e[:8] = 0.0
    chi_g[8:16] = 0.0
    return k4, p4, chi_g, chi_e

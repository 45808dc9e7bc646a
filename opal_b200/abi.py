"""ctypes binding of the C ABI in include/opal_b200.h.

`OpalABI(cdll, prefix)` binds every entry point the header declares; `Slab` is one context
(= one MPI rank of the reference = one x-slab on one GPU) with numpy in/out.  The product
library is bound with prefix ``opal_`` (see `opal_b200.lib`); the test-suite binds the CPU
oracle (its own prefix) through the same class, so parity tests drive both through
identical code.  Nothing in this module knows about the oracle.
"""
from __future__ import annotations

import ctypes as C
import numpy as np

OK, EINVAL, ECUDA, ENCCL, ENOMEM, EASSERT_CFL, EASSERT_INDEX, ECHI_RANGE, ECAPACITY, EASSERT_PROB = range(10)
STATUS_NAMES = ["OK", "EINVAL", "ECUDA", "ENCCL", "ENOMEM", "EASSERT_CFL", "EASSERT_INDEX", "ECHI_RANGE", "ECAPACITY", "EASSERT_PROB"]
BDY_INTERNAL, BDY_LASER, BDY_ABSORBING, BDY_CONDUCTING = range(4)
ELECTRON, ION, PHOTON = range(3)
LEFT, RIGHT = 0, 1
FLAG_NO_RADIATION_REACTION = 1 << 0
FLAG_NO_BEAMING = 1 << 1
FLAG_IMMOBILE_PHOTONS = 1 << 2
FLAG_NO_STIMULATED_EMISSION = 1 << 3
FLAG_PAIR_CREATION = 1 << 4      # extension: nonlinear Breit-Wheeler pair creation (no reference counterpart)
FLAG_LEAN_STATE = 1 << 8
FLAG_STRICT_FP = 1 << 9
GHOST_SIZE, LASER_BDY_SIZE, ABSORBING_BDY_SIZE, CONDUCTING_BDY_SIZE = 4, 4, 200, 4
CELL_DOUBLES, HALO_CELLS = 11, 8

MIGRANT_DTYPE = np.dtype([
    ("cell", "<i8"), ("x", "<f8"), ("prev_x", "<f8"), ("u", "<f8", (3,)), ("weight", "<f8"), ("gamma", "<f8"),
    ("chi", "<f8"), ("tau", "<f8"), ("work", "<f8"), ("y", "<f8"), ("z", "<f8"), ("birth_time", "<f8"),
    ("pad", "<f8", (2,))])
assert MIGRANT_DTYPE.itemsize == 128

_i64, _f64, _vp, _int, _u64, _u32 = C.c_int64, C.c_double, C.c_void_p, C.c_int, C.c_uint64, C.c_uint32
_pi64, _pf64 = C.POINTER(C.c_int64), C.POINTER(C.c_double)

# name -> argtypes; every function returns opal_status (int) except last_error.  This table IS
# the header: tests/test_abi_symbols.py checks it against include/opal_b200.h.
SIGNATURES = {
    "create": [C.POINTER(_vp), _int, _i64, _f64, _f64, _f64, _int, _int, _int, _int, _u64, _u32],
    "destroy": [_vp],
    "last_error": [_vp],
    "grid_shape": [_vp, _pi64, _pi64, _pi64],
    "particles_upload": [_vp, _int, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "set_ion": [_vp, _f64, _f64],
    "particles_count": [_vp, _int, _pi64],
    "particles_download": [_vp, _int, _i64] + [_vp] * 14,
    "fields_upload": [_vp, _vp],
    "fields_download": [_vp, _vp],
    "advance": [_vp, _int],
    "last_leavers": [_vp, _int, _pi64, _pi64],
    "migrants_take": [_vp, _int, _int, _i64, _pi64, _vp],
    "migrants_give": [_vp, _int, _int, _i64, _vp],
    "emit_radiation": [_vp, _f64, _pf64, _pf64, _pf64],
    "absorb": [_vp, _f64, _pf64, _pf64],
    "set_absorption": [_vp, _int, _pf64, _pf64],
    "photon_pair_depths": [_vp, _int, _i64, _vp, _vp],
    "clear": [_vp],
    "deposit": [_vp, _int],
    "halo_take": [_vp, _int, _vp],
    "halo_give": [_vp, _int, _vp],
    "synchronize": [_vp, _f64, _f64, _f64],
    "laser_x": [_vp, _pf64],
    "field_advance": [_vp],
    "run_steps": [_vp, _i64, _f64, _vp, _vp, _int, _int, _vp],
    "energy": [_vp, _pf64, _pf64, _pf64, _pf64],
    "hgram_minmax": [_vp, _int, _int, _vp, _vp, _vp, _vp, _pi64],
    "hgram_bin": [_vp, _int, _int, _vp, _vp, _vp, _vp, _vp, _int, _int, _vp, _pf64],
    "hgram_bin_static": [_vp, _int, _int, _vp, _vp, _vp, _vp, _vp, _int, _int, _vp, _pf64],
    "comm_unique_id": [_vp],
    "comm_init": [_vp, _vp],
    "stream": [_vp, C.POINTER(_vp)],
    "launch_count": [_vp, _pi64],
    "timing": [_vp, _int],
    "kernel_times": [_vp, _vp, _vp],
    "burst_steps": [_vp, _pi64],
    "particles_generate": [_vp, _int, _i64, _f64, _vp, _vp, _int, _i64, _u64],
    "qed_rate": [_i64, _vp, _vp, _vp],
    "qed_sample": [_i64, _vp, _vp, _vp, _vp, _vp, _int, _vp],
    "qed_airy": [_i64, _vp, _vp, _vp],
    "qed_pair_cross_sections": [_i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "fields_at": [_vp, _i64, _vp, _vp, _vp],
    "bw_rate": [_i64, _vp, _vp],
    "bw_sample": [_i64, _vp, _vp, _vp, _vp],
    "pairs_created": [_vp, _pi64],
}


class OpalError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str = ""):
        self.status = status
        name = STATUS_NAMES[status] if 0 <= status < len(STATUS_NAMES) else str(status)
        super().__init__(f"{where}: OPAL_{name}" + (f" ({detail})" if detail else ""))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64arr(a, n=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if n is not None and a.shape != (n,):
        raise ValueError(f"expected shape ({n},), got {a.shape}")
    return a


class OpalABI:
    """All entry points of include/opal_b200.h bound on `cdll` under `prefix`."""

    def __init__(self, cdll: C.CDLL, prefix: str = "opal_"):
        self.cdll, self.prefix = cdll, prefix
        for name, argtypes in SIGNATURES.items():
            fn = getattr(cdll, prefix + name)   # AttributeError == missing export: fail loudly
            fn.argtypes = argtypes
            fn.restype = C.c_char_p if name == "last_error" else C.c_int
            setattr(self, name, fn)

    # -- pure functions of the path -------------------------------------------------------
    def rate(self, chi, gamma):
        chi, gamma = _f64arr(np.atleast_1d(chi)), _f64arr(np.atleast_1d(gamma))
        out = np.empty_like(chi)
        st = self.qed_rate(chi.size, _ptr(chi), _ptr(gamma), _ptr(out))
        if st:
            raise OpalError(st, "qed_rate")
        return out

    def sample(self, chi, gamma, r1, r2, r3, classical=False):
        chi = _f64arr(np.atleast_1d(chi))
        n = chi.size
        gamma, r1, r2, r3 = (_f64arr(np.broadcast_to(np.atleast_1d(v), (n,))) for v in (gamma, r1, r2, r3))
        out = np.empty((n, 3))
        st = self.qed_sample(n, _ptr(chi), _ptr(gamma), _ptr(r1), _ptr(r2), _ptr(r3), int(classical), _ptr(out))
        if st:
            raise OpalError(st, "qed_sample")
        return out

    def bw_f(self, chi):
        """F(chi) = chi T(chi) of the Breit-Wheeler extension (rate = alpha_f F / (tau_C k0))"""
        chi = _f64arr(np.atleast_1d(chi))
        out = np.empty_like(chi)
        st = self.bw_rate(chi.size, _ptr(chi), _ptr(out))
        if st:
            raise OpalError(st, "bw_rate")
        return out

    def bw_share(self, chi, r1, r2):
        """the electron's share f of the photon energy in a Breit-Wheeler pair, for uniforms r1, r2"""
        chi = _f64arr(np.atleast_1d(chi))
        n = chi.size
        r1, r2 = (_f64arr(np.broadcast_to(np.atleast_1d(v), (n,))) for v in (r1, r2))
        out = np.empty(n)
        st = self.bw_sample(n, _ptr(chi), _ptr(r1), _ptr(r2), _ptr(out))
        if st:
            raise OpalError(st, "bw_sample")
        return out

    def airy(self, x):
        """(Ai(x), defined) -- airy_ai_for_positive (airy.rs:19); defined is False where the reference gives None"""
        x = _f64arr(np.atleast_1d(x))
        ai, ok = np.empty_like(x), np.empty(x.size, dtype=np.int32)
        st = self.qed_airy(x.size, _ptr(x), _ptr(ai), _ptr(ok))
        if st:
            raise OpalError(st, "qed_airy")
        return ai, ok.astype(bool)

    def pair_cross_sections(self, k4, p4, chi_gamma, chi_e):
        """(sigma_abs, sigma_st, abs_defined, st_defined) for n pairs (photon_absorption.rs:17, stimulated_emission.rs:17)"""
        k4, p4 = _f64arr(np.atleast_2d(k4)), _f64arr(np.atleast_2d(p4))
        n = k4.shape[0]
        chi_gamma, chi_e = (_f64arr(np.broadcast_to(np.atleast_1d(v), (n,))) for v in (chi_gamma, chi_e))
        sa, ss, ok = np.empty(n), np.empty(n), np.empty(n, dtype=np.int32)
        st = self.qed_pair_cross_sections(n, _ptr(k4), _ptr(p4), _ptr(chi_gamma), _ptr(chi_e), _ptr(sa), _ptr(ss), _ptr(ok))
        if st:
            raise OpalError(st, "qed_pair_cross_sections")
        return sa, ss, (ok & 1).astype(bool), (ok & 2).astype(bool)


# opal_quantity (particle/mod.rs:455-470) and their units (mod.rs:475-486)
QUANTITIES = ("x", "r", "energy", "px", "py", "pz", "p_perp", "theta", "phi", "longitude", "latitude", "work", "chi")
QUANTITY_UNITS = {"x": "m", "r": "m", "energy": "MeV", "px": "MeV/c", "py": "MeV/c", "pz": "MeV/c", "p_perp": "MeV/c", "theta": "rad",
                  "phi": "rad", "longitude": "rad", "latitude": "rad", "work": "J", "chi": "1"}
W_WEIGHT, W_ENERGY = 0, 1
HGRAM_STATIC_MAX_BINS = 65536      # OPAL_HGRAM_STATIC_MAX_BINS

PARTICLE_FIELDS = ("cell", "x", "prev_x", "ux", "uy", "uz", "weight", "gamma", "chi", "tau", "work", "y", "z", "birth_time")


class Slab:
    """One context: an x-slab of the Yee grid and the particles inside it (one reference rank)."""

    def __init__(self, abi: OpalABI, nx, xmin, dx, dt=None, left=BDY_INTERNAL, right=BDY_INTERNAL, rank=0, nranks=1,
                 seed=0, flags=0, device=0):
        from .constants import SPEED_OF_LIGHT
        self.abi = abi
        self.nx, self.xmin, self.dx = int(nx), float(xmin), float(dx)
        self.dt = 0.95 * self.dx / SPEED_OF_LIGHT if dt is None else float(dt)   # main.rs:71
        self.rank, self.nranks, self.flags = rank, nranks, flags
        h = C.c_void_p()
        st = abi.create(C.byref(h), device, self.nx, self.xmin, self.dx, self.dt, left, right, rank, nranks, seed, flags)
        if st:
            raise OpalError(st, "create")
        self._h = h
        tot, l, r = C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(abi.grid_shape(h, C.byref(tot), C.byref(l), C.byref(r)), "grid_shape")
        self.ncells_total, self.left_size, self.right_size = tot.value, l.value, r.value

    def _ck(self, st, where):
        if st:
            msg = self.abi.last_error(self._h)
            raise OpalError(st, where, msg.decode() if msg else "")

    def close(self):
        if getattr(self, "_h", None):
            self.abi.destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- state ------------------------------------------------------------------------------
    def upload(self, species, cell, x, ux, uy, uz, weight, tau=None):
        cell = np.ascontiguousarray(cell, dtype=np.int64)
        n = cell.size
        x, ux, uy, uz = (_f64arr(v, n) for v in (x, ux, uy, uz))
        weight = _f64arr(np.broadcast_to(weight, (n,)))
        tau = None if tau is None else _f64arr(tau, n)
        self._ck(self.abi.particles_upload(self._h, species, n, _ptr(cell), _ptr(x), _ptr(ux), _ptr(uy), _ptr(uz),
                                           _ptr(weight), _ptr(tau)), "particles_upload")

    def set_ion(self, z_star, mass_number):
        self._ck(self.abi.set_ion(self._h, z_star, mass_number), "set_ion")

    def count(self, species):
        n = C.c_int64()
        self._ck(self.abi.particles_count(self._h, species, C.byref(n)), "particles_count")
        return n.value

    def download(self, species, fields=PARTICLE_FIELDS):
        n = self.count(species)
        out = {}
        ptrs = []
        for f in PARTICLE_FIELDS:
            if f in fields:
                out[f] = np.empty(n, dtype=np.int64 if f == "cell" else np.float64)
                ptrs.append(_ptr(out[f]))
            else:
                ptrs.append(None)
        self._ck(self.abi.particles_download(self._h, species, n, *ptrs), "particles_download")
        return out

    def fields(self):
        a = np.empty((self.ncells_total, CELL_DOUBLES))
        self._ck(self.abi.fields_download(self._h, _ptr(a)), "fields_download")
        return a

    def set_fields(self, cells11):
        a = _f64arr(cells11).reshape(self.ncells_total, CELL_DOUBLES)
        self._ck(self.abi.fields_upload(self._h, _ptr(a)), "fields_upload")

    # -- loop body --------------------------------------------------------------------------
    def advance(self, species):
        self._ck(self.abi.advance(self._h, species), "advance")

    def last_leavers(self, species):
        l, r = C.c_int64(), C.c_int64()
        self._ck(self.abi.last_leavers(self._h, species, C.byref(l), C.byref(r)), "last_leavers")
        return l.value, r.value

    def migrants_take(self, species, side, cap=1 << 16):
        while True:
            buf = np.zeros(cap, dtype=MIGRANT_DTYPE)
            n = C.c_int64()
            st = self.abi.migrants_take(self._h, species, side, cap, C.byref(n), _ptr(buf))
            if st == ECAPACITY and n.value > cap:
                cap = n.value
                continue
            self._ck(st, "migrants_take")
            return buf[: n.value].copy()

    def migrants_give(self, species, side, migrants):
        m = np.ascontiguousarray(migrants, dtype=MIGRANT_DTYPE)
        self._ck(self.abi.migrants_give(self._h, species, side, m.size, _ptr(m)), "migrants_give")

    def emit_radiation(self, t, min_energy_mev=None, max_angle=None, max_formation_length=None):
        def opt(v):
            return None if v is None else C.byref(C.c_double(v))
        self._ck(self.abi.emit_radiation(self._h, t, opt(min_energy_mev), opt(max_angle), opt(max_formation_length)),
                 "emit_radiation")

    def absorb(self, t, max_displacement=None, stop_time=None):
        """absorb (particle/interactions.rs:145)"""
        def opt(v):
            return None if v is None else C.byref(C.c_double(v))
        self._ck(self.abi.absorb(self._h, t, opt(max_displacement), opt(stop_time)), "absorb")

    def set_absorption(self, enabled, max_displacement=None, stop_time=None):
        """run_steps then calls absorb every step (main.rs:246-248)"""
        def opt(v):
            return None if v is None else C.byref(C.c_double(v))
        self._ck(self.abi.set_absorption(self._h, int(enabled), opt(max_displacement), opt(stop_time)), "set_absorption")

    def photon_pair_depths(self):
        """(tau_abs, tau_st) of the photons, in download order (photon.rs:28-29)"""
        n = self.count(PHOTON)
        a, b = np.empty(n), np.empty(n)
        self._ck(self.abi.photon_pair_depths(self._h, 0, n, _ptr(a), _ptr(b)), "photon_pair_depths")
        return a, b

    def set_photon_pair_depths(self, tau_abs, tau_st):
        a, b = _f64arr(tau_abs).copy(), _f64arr(tau_st).copy()
        self._ck(self.abi.photon_pair_depths(self._h, 1, a.size, _ptr(a), _ptr(b)), "photon_pair_depths")

    def clear(self):
        self._ck(self.abi.clear(self._h), "clear")

    def deposit(self, species):
        self._ck(self.abi.deposit(self._h, species), "deposit")

    def halo_take(self, side):
        a = np.empty((HALO_CELLS, CELL_DOUBLES))
        self._ck(self.abi.halo_take(self._h, side, _ptr(a)), "halo_take")
        return a

    def halo_give(self, side, cells):
        a = _f64arr(cells).reshape(HALO_CELLS, CELL_DOUBLES)
        self._ck(self.abi.halo_give(self._h, side, _ptr(a)), "halo_give")

    def synchronize(self, t, laser_y=0.0, laser_z=0.0):
        self._ck(self.abi.synchronize(self._h, t, laser_y, laser_z), "synchronize")

    def laser_x(self):
        x = C.c_double()
        self._ck(self.abi.laser_x(self._h, C.byref(x)), "laser_x")
        return x.value

    def field_advance(self):
        self._ck(self.abi.field_advance(self._h), "field_advance")

    def run_steps(self, nsteps, t0=0.0, laser_y=None, laser_z=None, current_deposition=True, photon_emission=False,
                  emit_filters=None):
        ly = None if laser_y is None else _f64arr(laser_y, nsteps)
        lz = None if laser_z is None else _f64arr(laser_z, nsteps)
        f3 = None
        if emit_filters is not None:
            f3 = np.array([np.nan if v is None else v for v in emit_filters], dtype=np.float64)
        self._ck(self.abi.run_steps(self._h, nsteps, t0, _ptr(ly), _ptr(lz), int(current_deposition),
                                    int(photon_emission), _ptr(f3)), "run_steps")

    def energy(self):
        v = [C.c_double() for _ in range(4)]
        self._ck(self.abi.energy(self._h, *[C.byref(x) for x in v]), "energy")
        return tuple(x.value for x in v)   # (em, electrons, ions, photons), joules, this slab

    def hgram_minmax(self, species, quantities, log_scaled=None):
        """pass 1 of Histogram::generate_* (hgram.rs:186-204): (min[], max[], n) of this slab"""
        nd = len(quantities)
        q = np.array([QUANTITIES.index(v) if isinstance(v, str) else int(v) for v in quantities], dtype=np.int32)
        lg = np.array([int(bool(v)) for v in (log_scaled or [0] * nd)], dtype=np.int32)
        mn, mx, n = np.zeros(2), np.zeros(2), C.c_int64(0)
        self._ck(self.abi.hgram_minmax(self._h, species, nd, _ptr(q), _ptr(lg), _ptr(mn), _ptr(mx), C.byref(n)), "hgram_minmax")
        return mn[:nd].copy(), mx[:nd].copy(), n.value

    def hgram_bin(self, species, quantities, gmin, bin_sz, nbins, log_scaled=None, weight=W_WEIGHT, log_density_adjust=True, static=False):
        """pass 2 (hgram.rs:222-258 / 333-366): (weighted counts of this slab before /bin volume, flat; total weight).
        static=True goes through opal_hgram_bin_static (no device allocation; at most HGRAM_STATIC_MAX_BINS bins)"""
        nd = len(quantities)
        q = np.array([QUANTITIES.index(v) if isinstance(v, str) else int(v) for v in quantities], dtype=np.int32)
        lg = np.array([int(bool(v)) for v in (log_scaled or [0] * nd)], dtype=np.int32)
        gmin, bin_sz = _f64arr(np.atleast_1d(gmin), nd), _f64arr(np.atleast_1d(bin_sz), nd)
        nb = np.ascontiguousarray(np.atleast_1d(nbins), dtype=np.int64)
        counts = np.zeros(int(np.prod(nb)))
        total = C.c_double(0.0)
        fn, what = (self.abi.hgram_bin_static, "hgram_bin_static") if static else (self.abi.hgram_bin, "hgram_bin")
        self._ck(fn(self._h, species, nd, _ptr(q), _ptr(lg), _ptr(gmin), _ptr(bin_sz), _ptr(nb), int(weight),
                    int(bool(log_density_adjust)), _ptr(counts), C.byref(total)), what)
        return counts, total.value

    def fields_at(self, cell, xi):
        cell = np.ascontiguousarray(np.atleast_1d(cell), dtype=np.int64)
        xi = _f64arr(np.atleast_1d(xi), cell.size)
        out = np.empty((cell.size, 6))
        self._ck(self.abi.fields_at(self._h, cell.size, _ptr(cell), _ptr(xi), _ptr(out)), "fields_at")
        return out[:, :3], out[:, 3:]

    # -- multi-GPU / instrumentation --------------------------------------------------------
    def comm_init(self, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, 128)
        self._ck(self.abi.comm_init(self._h, buf), "comm_init")

    def stream(self):
        s = C.c_void_p()
        self._ck(self.abi.stream(self._h, C.byref(s)), "stream")
        return s.value or 0

    def pairs_created(self):
        n = C.c_int64(0)
        self._ck(self.abi.pairs_created(self._h, C.byref(n)), "pairs_created")
        return n.value

    def launch_count(self):
        n = C.c_int64()
        self._ck(self.abi.launch_count(self._h, C.byref(n)), "launch_count")
        return n.value

    def burst_steps(self):
        n = C.c_int64(0)
        self._ck(self.abi.burst_steps(self._h, C.byref(n)), "burst_steps")
        return n.value

    def timing(self, enable=True):
        self._ck(self.abi.timing(self._h, int(enable)), "timing")

    def kernel_times(self):
        ms, n = np.zeros(3), np.zeros(3, dtype=np.int64)
        self._ck(self.abi.kernel_times(self._h, _ptr(ms), _ptr(n)), "kernel_times")
        return {"push": (ms[0], int(n[0])), "scatter": (ms[1], int(n[1])), "field": (ms[2], int(n[2]))}

    def generate(self, species, npc, weight, u_th=(0.0, 0.0, 0.0), drift=(0.0, 0.0, 0.0), two_stream=False,
                 global_cell_offset=0, seed=12345):
        a, b = _f64arr(u_th, 3), _f64arr(drift, 3)
        self._ck(self.abi.particles_generate(self._h, species, npc, weight, _ptr(a), _ptr(b), int(two_stream),
                                             global_cell_offset, seed), "particles_generate")

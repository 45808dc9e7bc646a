"""An in-process ring of CPU-oracle slabs: the reference's MPI decomposition (one YeeGrid + Populations per rank,
grid/mod.rs:122-151) run rank after rank in ONE process, leavers and halo strips handed between neighbours through the
host-staged calls in the order of mod.rs:290-373 / yee.rs:351-452.  It is the checker for the multi-GPU path: slab g of
this ring is what GPU g must reproduce -- per step and per direction the leaver counts (`Population::advance` returns
`(gone_left, gone_right)`, mod.rs:284) bit for bit, particles and fields to 1e-10.

Test infrastructure (loads oracle/libopal_oracle.so): used by tests/mgpu_check.py, tests/test_multislab_gloo.py and the
`multi_gpu_parity` block of bench.py."""
import ctypes
import pathlib

import numpy as np

from opal_b200.abi import BDY_INTERNAL, ELECTRON, ION, LEFT, PHOTON, RIGHT, OpalABI, Slab
from opal_b200.constants import SPEED_OF_LIGHT

ROOT = pathlib.Path(__file__).resolve().parent.parent


def load_oracle():
    return OpalABI(ctypes.CDLL(str(ROOT / "oracle" / "libopal_oracle.so")), "opal_ref_")


class OracleRing:
    def __init__(self, oracle, design, flags=0):
        self.design, self.n = design, design.numtasks
        dt = 0.95 * design.dx / SPEED_OF_LIGHT
        self.slabs = [Slab(oracle, design.nx[r], design.slab_xmin(r), design.dx, dt, design.left, design.right, rank=r, nranks=self.n,
                           seed=r, flags=flags) for r in range(self.n)]
        self.dt = dt
        self.offsets = [sum(design.nx[:r]) for r in range(self.n)]
        self.has_left = [not (r == 0 and design.left != BDY_INTERNAL) for r in range(self.n)]
        self.has_right = [not (r == self.n - 1 and design.right != BDY_INTERNAL) for r in range(self.n)]

    def close(self):
        for s in self.slabs:
            s.close()

    def upload(self, species, p):
        """p: global-cell particle dict; every slab takes its share (an empty share still registers the species)"""
        for r, s in enumerate(self.slabs):
            off = self.offsets[r]
            mine = (p["cell"] >= off) & (p["cell"] < off + self.design.nx[r])
            s.upload(species, p["cell"][mine] - off, p["x"][mine], p["ux"][mine], p["uy"][mine], p["uz"][mine], p["weight"][mine])

    def advance(self, species):
        """Population::advance on every rank, then the neighbour exchange; returns [(gone_left, gone_right)] per slab"""
        n = self.n
        for s in self.slabs:
            s.advance(species)
        gone = [s.last_leavers(species) for s in self.slabs]
        if n > 1:
            out_l = [s.migrants_take(species, LEFT) if self.has_left[r] else None for r, s in enumerate(self.slabs)]
            out_r = [s.migrants_take(species, RIGHT) if self.has_right[r] else None for r, s in enumerate(self.slabs)]
            for r, s in enumerate(self.slabs):
                if self.has_left[r]:
                    m = out_r[(r - 1) % n]           # what my left neighbour sent to its right
                    if m is not None and m.size:
                        s.migrants_give(species, LEFT, m)
                if self.has_right[r]:
                    m = out_l[(r + 1) % n]
                    if m is not None and m.size:
                        s.migrants_give(species, RIGHT, m)
        return gone

    def synchronize(self, t, laser_y=0.0, laser_z=0.0):
        n = self.n
        if n > 1:
            out_l = [s.halo_take(LEFT) if self.has_left[r] else None for r, s in enumerate(self.slabs)]
            out_r = [s.halo_take(RIGHT) if self.has_right[r] else None for r, s in enumerate(self.slabs)]
            for r, s in enumerate(self.slabs):
                if self.has_left[r]:
                    s.halo_give(LEFT, out_r[(r - 1) % n])
                if self.has_right[r]:
                    s.halo_give(RIGHT, out_l[(r + 1) % n])
        for r, s in enumerate(self.slabs):
            s.synchronize(t, laser_y if r == 0 else 0.0, laser_z if r == 0 else 0.0)

    def step(self, t, laser=None, species=(ELECTRON,), current_deposition=True):
        """one iteration of main.rs:238-267 on every rank; returns (t + dt, {species: [(gone_l, gone_r)] per slab})"""
        gone = {sp: self.advance(sp) for sp in species}
        if current_deposition:
            for s in self.slabs:
                s.clear()
                for sp in species:
                    if sp != PHOTON:
                        s.deposit(sp)
        if laser is not None:
            x2 = self.slabs[0].laser_x()
            self.synchronize(t, laser[0](t, x2), laser[1](t, x2))
        else:
            self.synchronize(t)
        for s in self.slabs:
            s.field_advance()
        return t + self.dt, gone

    def download(self, species, fields):
        """per-slab particle dicts with GLOBAL cell indices"""
        out = []
        for r, s in enumerate(self.slabs):
            d = s.download(species, fields=fields)
            d["cell"] = d["cell"] + self.offsets[r]
            out.append(d)
        return out

    def interior_fields(self):
        return [s.fields()[s.left_size: s.left_size + self.design.nx[r]] for r, s in enumerate(self.slabs)]


def density_of(p, nx):
    """ne(x) per cell from a particle set: what GridDesign.balanced integrates (grid/mod.rs:157-206)"""
    return np.bincount(p["cell"], weights=p["weight"], minlength=nx).astype(float)

"""Host-side domain decomposition: the reference's GridDesign (src/grid/mod.rs:122-206) and the
ring-of-slabs time loop (src/main.rs:238-267 with np > 1).

`GridDesign.unbalanced / balanced` reproduce the reference's slab sizes and offsets (including its
quirks: `unbalanced` silently drops `nx mod np` cells, mod.rs:128; `balanced` reserves `min_size`
cells for the last rank, mod.rs:164).

`SlabRing` drives one slab per rank.  Two exchange modes:
  * "nccl":  the library owns an NCCL communicator (opal_comm_init) and exchanges halos and migrants
             itself inside opal_run_steps / opal_advance / opal_synchronize -- the production path;
  * "host":  halos and migrants are taken from / given to the context through host buffers and moved
             with torch.distributed point-to-point calls in the reference's even/odd order
             (yee.rs:363-409, mod.rs:308-354).  Works with any backend (gloo on CPU).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Callable, List, Optional

import numpy as np

from .abi import (BDY_INTERNAL, CELL_DOUBLES, ELECTRON, HALO_CELLS, ION, LEFT, MIGRANT_DTYPE, PHOTON, RIGHT, OpalABI, Slab)
from .constants import SPEED_OF_LIGHT

MIN_SIZE = 8      # YeeGrid::min_size() = 2 * GHOST_SIZE  (yee.rs:783-785)


@dataclass
class GridDesign:
    """grid/mod.rs:104-120"""
    numtasks: int
    left: int
    right: int
    xmin: float
    dx: float
    nx: List[int]
    offset: List[float]

    @staticmethod
    def _offsets(nx, dx):
        out, rt = [], 0
        for n in nx:
            out.append(float(rt) * dx)     # (*rt as f64) * dx  mod.rs:134
            rt += n
        return out

    @classmethod
    def unbalanced(cls, numtasks, size, xmin, dx, min_subsize=MIN_SIZE, left=BDY_INTERNAL, right=BDY_INTERNAL):
        """mod.rs:126-151"""
        subsize = max(size // numtasks, min_subsize)
        nx = [subsize] * numtasks
        return cls(numtasks, left, right, xmin, dx, nx, cls._offsets(nx, dx))

    @classmethod
    def balanced(cls, numtasks, size, xmin, dx, ne: Callable[[float], float], min_subsize=MIN_SIZE, left=BDY_INTERNAL,
                 right=BDY_INTERNAL):
        """mod.rs:157-206: equal integral of ne dx per rank, at least min_subsize cells each"""
        if numtasks == 0:
            return cls.unbalanced(numtasks, size, xmin, dx, min_subsize, left, right)
        ppc = [dx * ne(xmin + float(n) * dx) for n in range(size - min_subsize)]
        cumsum = np.cumsum(np.asarray(ppc, dtype=np.float64)) if ppc else np.zeros(0)
        # sequential f64 accumulation like the reference's scan (np.cumsum is sequential for float64)
        target = cumsum[-1] / float(numtasks)
        ncells, start = [], 0
        for p in range(1, numtasks):
            tail = cumsum[start + min_subsize:]
            hits = np.nonzero(tail >= target * float(p))[0]
            if hits.size == 0:
                raise ValueError("balanced: cannot place partition (the reference's unwrap() would panic, mod.rs:179)")
            i = int(hits[0])
            ncells.append(i + min_subsize)
            start += i + min_subsize
        ncells.append(size - sum(ncells))
        return cls(numtasks, left, right, xmin, dx, ncells, cls._offsets(ncells, dx))

    def slab_xmin(self, rank):
        return self.xmin + self.offset[rank]          # yee.rs:296


class SlabRing:
    """This rank's slab of a ring of `nranks` slabs + the per-step exchange with its neighbours."""

    def __init__(self, abi: OpalABI, design: GridDesign, rank: int, dt: Optional[float] = None, flags=0, seed=None, device=0,
                 exchange="host", dist=None):
        self.abi, self.design, self.rank, self.nranks = abi, design, rank, design.numtasks
        self.exchange, self.dist = exchange, dist
        dt = 0.95 * design.dx / SPEED_OF_LIGHT if dt is None else dt       # main.rs:71
        self.slab = Slab(abi, design.nx[rank], design.slab_xmin(rank), design.dx, dt, design.left, design.right, rank=rank,
                         nranks=self.nranks, seed=rank if seed is None else seed, flags=flags, device=device)   # main.rs:53
        self.left_rank = (rank - 1 + self.nranks) % self.nranks            # yee.rs:309
        self.right_rank = (rank + 1) % self.nranks                         # yee.rs:310
        # yee.rs:329-343: a side has a neighbour iff its boundary is Internal
        self.has_left = not (rank == 0 and design.left != BDY_INTERNAL)
        self.has_right = not (rank == self.nranks - 1 and design.right != BDY_INTERNAL)
        if exchange == "nccl" and self.nranks > 1:
            self._init_nccl()

    # -- NCCL bootstrap: rank 0 creates the unique id, torch.distributed broadcasts it ----------
    def _init_nccl(self):
        import torch
        buf = ctypes.create_string_buffer(128)
        if self.rank == 0:
            st = self.abi.comm_unique_id(buf)
            if st:
                raise RuntimeError(f"opal_comm_unique_id failed: {st}")
        dev = "cuda" if self.dist.get_backend() == "nccl" else "cpu"
        t = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).to(dev)
        self.dist.broadcast(t, 0)
        self.slab.comm_init(bytes(t.cpu().numpy().tobytes()))

    # -- host-staged neighbour exchange in the reference's even/odd order ----------------------
    def _sendrecv(self, to_left: Optional[np.ndarray], to_right: Optional[np.ndarray], nbytes_fixed: Optional[int]):
        """returns (from_left, from_right) as uint8 arrays (None where there is no neighbour).
        nbytes_fixed None => variable-length messages (a length word is exchanged first)."""
        import torch
        dist = self.dist

        def send(arr, peer):
            a = np.frombuffer(arr.tobytes(), dtype=np.uint8)
            if nbytes_fixed is None:
                dist.send(torch.tensor([a.size], dtype=torch.int64), peer)
            if a.size:
                dist.send(torch.from_numpy(a.copy()), peer)

        def recv(peer):
            if nbytes_fixed is None:
                n = torch.zeros(1, dtype=torch.int64)
                dist.recv(n, peer)
                n = int(n.item())
            else:
                n = nbytes_fixed
            t = torch.zeros(n, dtype=torch.uint8)
            if n:
                dist.recv(t, peer)
            return t.numpy()

        from_left = from_right = None
        even = self.rank % 2 == 0
        # phase 1: even ranks send right then receive from the right; odd ranks receive from the left then send left
        # phase 2: mirrored (yee.rs:363-409, mod.rs:308-354)
        for phase in (0, 1):
            right_side = (phase == 0) == even
            if right_side:
                if self.has_right:
                    if even:
                        send(to_right, self.right_rank); from_right = recv(self.right_rank)
                    else:
                        from_right = recv(self.right_rank); send(to_right, self.right_rank)
            else:
                if self.has_left:
                    if even:
                        send(to_left, self.left_rank); from_left = recv(self.left_rank)
                    else:
                        from_left = recv(self.left_rank); send(to_left, self.left_rank)
        return from_left, from_right

    def advance(self, species):
        s = self.slab
        s.advance(species)                               # with exchange == "nccl" this already migrated
        if self.exchange == "host" and self.nranks > 1:
            out_l = s.migrants_take(species, LEFT) if self.has_left else None
            out_r = s.migrants_take(species, RIGHT) if self.has_right else None
            in_l, in_r = self._sendrecv(out_l, out_r, None)
            if in_l is not None and in_l.size:
                s.migrants_give(species, LEFT, np.frombuffer(in_l.tobytes(), dtype=MIGRANT_DTYPE))
            if in_r is not None and in_r.size:
                s.migrants_give(species, RIGHT, np.frombuffer(in_r.tobytes(), dtype=MIGRANT_DTYPE))

    def synchronize(self, t, laser_y=0.0, laser_z=0.0):
        s = self.slab
        if self.exchange == "host" and self.nranks > 1:
            out_l = s.halo_take(LEFT) if self.has_left else None
            out_r = s.halo_take(RIGHT) if self.has_right else None
            nb = HALO_CELLS * CELL_DOUBLES * 8
            in_l, in_r = self._sendrecv(out_l, out_r, nb)
            if in_l is not None:
                s.halo_give(LEFT, np.frombuffer(in_l.tobytes(), dtype=np.float64))
            if in_r is not None:
                s.halo_give(RIGHT, np.frombuffer(in_r.tobytes(), dtype=np.float64))
        s.synchronize(t, laser_y, laser_z)

    def step(self, t, laser=None, current_deposition=True, emission=None, species=(ELECTRON, ION, PHOTON)):
        """one iteration of main.rs:238-267 call by call"""
        s = self.slab
        for sp in species:
            self.advance(sp)
        if emission is not None:
            s.emit_radiation(t, *emission)
        if current_deposition:
            s.clear()
            for sp in species:
                if sp != PHOTON:
                    s.deposit(sp)
        if laser is not None and self.rank == 0:
            x2 = s.laser_x()
            self.synchronize(t, laser[0](t, x2), laser[1](t, x2))
        else:
            self.synchronize(t, 0.0, 0.0)
        s.field_advance()
        return t + s.dt

    def energy(self):
        """write_energies (main.rs:23-42): local contributions summed over ranks (reduce_into_root)"""
        import torch
        e = torch.tensor(self.slab.energy(), dtype=torch.float64)
        if self.nranks > 1:
            if self.dist.get_backend() == "nccl":
                e = e.cuda()
            self.dist.all_reduce(e)
        return tuple(e.cpu().tolist())

"""Workload for the ncu capture of the block-per-cell launch (k_push_lean_fast<.., HEAVY>): one cold pile of 32 000 electrons
among 2048 cells of 100 (tools/bench_heavy_cell.py's case), a few steps."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
os.environ["OPAL_HEAVY_MIN"] = "1024"; os.environ["OPAL_HEAVY_FACTOR"] = "0"; os.environ["OPAL_HEAVY_FORCE"] = "1"; os.environ["OPAL_BURST"] = "0"
from opal_b200 import lib
from opal_b200.abi import ELECTRON, FLAG_LEAN_STATE, Slab
import pic_cases as pc
nx, dx, n_big = 2048, 500.0, 32000
a = Slab(lib.load(), nx, 0.0, dx, flags=FLAG_LEAN_STATE)
p1 = pc.thermal_particles(nx, n_big, dx, 1.0e-5, seed=1, cells=np.array([nx // 2]), ne=20.0 * n_big / 100)
p2 = pc.thermal_particles(nx, 100, dx, 0.02, seed=2, cells=np.setdiff1d(np.arange(nx), [nx // 2]))
p2["weight"] *= 1.000123
p = {k: np.concatenate([p1[k], p2[k]]) for k in p1}
p["weight"] *= 1.0e-12
pc.upload(a, ELECTRON, p)
a.run_steps(8)
print("particles", a.count(ELECTRON), "launches", a.launch_count())

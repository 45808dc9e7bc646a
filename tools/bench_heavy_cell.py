"""dev helper: what the block-per-cell launch buys for ONE very heavy cell (the lean strip kernel gives a cell to one warp):
a thin plasma with a single cell of N particles, steps per second with OPAL_HEAVY_MIN=0 (never) and with the threshold below N"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from opal_b200 import lib
from opal_b200.abi import ELECTRON, FLAG_LEAN_STATE, Slab
import pic_cases as pc

gpu = lib.load()
nx, dx = 2048, 500.0
for n_big in (1000, 2000, 4000, 8000, 16000, 32000):
    row = []
    for hm in ("0", "1024"):    # (the library's default threshold is 4096)
        os.environ["OPAL_HEAVY_MIN"] = hm
        os.environ["OPAL_HEAVY_FACTOR"] = "0"
        a = Slab(gpu, nx, 0.0, dx, flags=FLAG_LEAN_STATE)
        p1 = pc.thermal_particles(nx, n_big, dx, 1.0e-5, seed=1, cells=np.array([nx // 2]), ne=20.0 * n_big / 100)
        p2 = pc.thermal_particles(nx, 100, dx, 0.02, seed=2, cells=np.setdiff1d(np.arange(nx), [nx // 2]))
        p2["weight"] *= 1.000123
        p = {k: np.concatenate([p1[k], p2[k]]) for k in p1}
        p["weight"] *= 1.0e-12           # (no space charge to speak of: the pile stays where it is)
        pc.upload(a, ELECTRON, p)
        a.run_steps(300)                 # warm-up: the launch switches itself on, strips settle
        t0 = time.perf_counter()
        a.run_steps(2000)
        dt = time.perf_counter() - t0
        row.append(1e6 * dt / 2000)
        cnt = np.bincount(a.download(ELECTRON)["cell"], minlength=nx).max()
        a.close()
    print(f"one cell of {n_big} among {nx} cells of 100: {row[0]:.1f} us/step one warp per cell, {row[1]:.1f} us/step with the block-per-cell launch (largest cell at the end: {cnt})", flush=True)

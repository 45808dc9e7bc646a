"""opal_hgram_bin_static: the entry point for callers that may not allocate device memory (no counterpart in the reference
-- include/opal_b200.h).  Its bin counters live in a static device array shared
by every call on a GPU, so the checks that matter are: same output as opal_hgram_bin (the existing entry point, which
allocates its counters on demand) on CALLS IN A ROW that differ in shape -- anything a call leaves in the array must not leak into
the next one -- and a clear refusal beyond the stated maximum.

This file sorts last on purpose (pytest -x at the end of a round stops at the first failure)."""
import numpy as np
import pytest

from opal_b200.abi import ECAPACITY, EINVAL, ELECTRON, FLAG_LEAN_STATE, HGRAM_STATIC_MAX_BINS, PHOTON, W_ENERGY, W_WEIGHT, OpalError, Slab
import pic_cases as pc


def _both(slab, species, quantities, nbins, weight=W_WEIGHT, log_scaled=None):
    """(counts, total) of the existing entry point and of the static one on the same grid"""
    mn, mx, n = slab.hgram_minmax(species, quantities, log_scaled)
    assert n > 0
    sz = [(b - a) / k for a, b, k in zip(mn, mx, nbins)]
    want = slab.hgram_bin(species, quantities, mn, sz, nbins, log_scaled, weight)
    got = slab.hgram_bin(species, quantities, mn, sz, nbins, log_scaled, weight, static=True)
    return want, got


def _same(want, got):
    (cw, tw), (cg, tg) = want, got
    assert cw.shape == cg.shape and cw.sum() > 0
    # same kernel, same bins: only the order of the floating-point atomics differs between two launches (test_gpu_hgram.py's bound)
    assert abs(tg - tw) <= 1e-12 * abs(tw)
    assert np.allclose(cg, cw, rtol=1e-11, atol=1e-13 * np.abs(cw).max())
    assert np.array_equal(cg == 0.0, cw == 0.0)          # a counter the last call left behind would show up in an empty bin


@pytest.mark.gpu
def test_static_scratch_histogram_matches_the_allocating_entry_point(gpu):
    nx, npc, dx = 96, 150, 500.0
    a = Slab(gpu, nx, -1.0e4, dx)
    p = pc.thermal_particles(nx, npc, dx, 0.3, seed=3)
    p["weight"] = p["weight"] * (1.0 + 0.5 * np.random.default_rng(4).random(p["weight"].size))
    pc.upload(a, ELECTRON, p)
    rng = np.random.default_rng(12)
    n = 3000
    k = rng.standard_normal((n, 3)) * 50.0
    a.upload(PHOTON, np.sort(rng.integers(0, nx, n)), rng.random(n), k[:, 0], k[:, 1], k[:, 2], np.full(n, 2.5))

    # calls in a row, each on a different shape: the largest grid the array holds first (every counter and the total's slot
    # get written), then smaller ones whose counters and total sit where the first call left non-zero values
    _same(*_both(a, ELECTRON, ["x", "px"], [256, 256]))                       # 65536 bins == the maximum
    _same(*_both(a, ELECTRON, ["energy"], [50]))
    _same(*_both(a, ELECTRON, ["energy"], [50]))                              # the very same call twice
    _same(*_both(a, PHOTON, ["x", "px"], [17, 9], weight=W_ENERGY))
    _same(*_both(a, ELECTRON, ["py", "pz"], [64, 33], weight=W_ENERGY))
    _same(*_both(a, ELECTRON, ["energy"], [40], log_scaled=[1]))

    # one bin too many: refused with a message, nothing computed, the context stays usable
    mn, mx, _ = a.hgram_minmax(ELECTRON, ["x", "px"])
    with pytest.raises(OpalError) as e:
        a.hgram_bin(ELECTRON, ["x", "px"], mn, [1.0, 1.0], [HGRAM_STATIC_MAX_BINS, 2], static=True)
    assert e.value.status == ECAPACITY and "exceed the static scratch of 65536" in str(e.value)
    with pytest.raises(OpalError) as e:
        a.hgram_bin(ELECTRON, ["x"], mn[:1], [1.0], [HGRAM_STATIC_MAX_BINS + 1], static=True)
    assert e.value.status == ECAPACITY
    _same(*_both(a, ELECTRON, ["x"], [HGRAM_STATIC_MAX_BINS]))

    # a species that is not there: zeros, like opal_hgram_bin
    c = Slab(gpu, nx, 0.0, dx)
    counts, total = c.hgram_bin(ELECTRON, ["x"], [0.0], [1.0], [8], static=True)
    assert total == 0.0 and not counts.any()
    a.close()
    c.close()


@pytest.mark.gpu
def test_static_scratch_histogram_on_the_lean_layout(gpu):
    """after steps on the lean layout the last step's movers wait in the strips: settling them may allocate, so the static entry
    point says so instead of doing it; after any reader it answers like opal_hgram_bin.  Two contexts share the array."""
    nx, npc, dx = 128, 64, 500.0
    a = Slab(gpu, nx, -1.0e4, dx, flags=FLAG_LEAN_STATE)
    b = Slab(gpu, nx, -1.0e4, dx)
    p = pc.thermal_particles(nx, npc, dx, 0.3, seed=5)
    pc.upload(a, ELECTRON, p)
    pc.upload(b, ELECTRON, p)
    _same(*_both(a, ELECTRON, ["x", "px"], [40, 40]))
    a.run_steps(5)
    with pytest.raises(OpalError) as e:
        a.hgram_bin(ELECTRON, ["x"], [-1.0e4], [dx], [nx], static=True)
    assert e.value.status == EINVAL and "pending" in str(e.value)
    assert a.count(ELECTRON) == nx * npc                                      # a reader: settles the strips
    _same(*_both(a, ELECTRON, ["x", "px"], [40, 40]))
    _same(*_both(b, ELECTRON, ["px"], [1000]))                                # another context, same array
    _same(*_both(a, ELECTRON, ["x"], [nx]))
    a.close()
    b.close()


def test_limit_is_the_same_in_header_binding_and_oracle(oracle):
    """no GPU needed: the header's constant, the Python table's and the CPU forwarder's refusal agree"""
    import pathlib
    import re
    header = (pathlib.Path(__file__).resolve().parent.parent / "include" / "opal_b200.h").read_text()
    assert int(re.search(r"#define OPAL_HGRAM_STATIC_MAX_BINS (\d+)", header).group(1)) == HGRAM_STATIC_MAX_BINS == 65536
    s = Slab(oracle, 16, 0.0, 1.0)
    pc.upload(s, ELECTRON, pc.thermal_particles(16, 4, 1.0, 0.1))
    with pytest.raises(OpalError) as e:
        s.hgram_bin(ELECTRON, ["x"], [0.0], [1.0], [HGRAM_STATIC_MAX_BINS + 1], static=True)
    assert e.value.status == ECAPACITY
    cs, ts = s.hgram_bin(ELECTRON, ["x"], [0.0], [1.0], [16], static=True)
    cw, tw = s.hgram_bin(ELECTRON, ["x"], [0.0], [1.0], [16])
    assert np.array_equal(cs, cw) and ts == tw
    s.close()

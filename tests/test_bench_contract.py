"""bench.py's reference arm (the CPU leg the driver times beside the CUDA path) keeps its JSON contract
and uses every host core it may run on, also when launched by torchrun (which exports OMP_NUM_THREADS=1)."""
import json
import os
import pathlib
import subprocess
import sys

ROOT = pathlib.Path(__file__).resolve().parent.parent


def _run(env_extra):
    env = dict(os.environ, **env_extra)
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1",
                          "--cpu-nx", "200", "--npc", "50"], capture_output=True, text=True, env=env, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout.strip()


def test_reference_arm_line_and_thread_count():
    line = json.loads(_run({"RANK": "0", "WORLD_SIZE": "2", "OMP_NUM_THREADS": "1"}).splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "particle-steps/s" and line["higher_is_better"] is True
    assert line["n_gpus"] == 2 and line["steps"] == 1 and line["warmup"] == 1
    assert line["value"] > 0 and line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == line["value"]
    assert cb["cores"] == len(os.sched_getaffinity(0))
    assert line["config"]["particles"] == 200 * 50


def test_reference_arm_other_ranks_do_no_work():
    assert _run({"RANK": "1", "WORLD_SIZE": "2"}) == ""


def test_product_arm_fails_loudly_without_a_gpu():
    """no CPU fallback: bench.py's own arm refuses to run when there is no CUDA device (and never falls back to the oracle)"""
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--steps", "1", "--warmup", "1", "--nx", "100", "--npc", "10"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode != 0 and "no CUDA device" in (out.stderr + out.stdout)
    assert out.stdout.strip() == ""          # no JSON line, hence no number


def test_parity_leg_of_the_bench_line(oracle, oracle_cdll):
    """bench.py's `parity` block (the CPU leg's arrays through the CUDA path, compared with the oracle's result), exercised
    on the CPU with the oracle standing in for the product library: identical code on both sides => exact agreement"""
    sys.path.insert(0, str(ROOT))
    import bench
    rate, secs, threads, n, state = bench.cpu_reference_run(200, 50, 2, 1, keep_state=True)
    assert n == 10000 and rate > 0 and state["steps"] == 3
    out = bench.gpu_parity_leg(oracle, state, 200, 0)
    assert out["ok"] and out["cells_bit_exact"] and out["n"] == 10000 and out["max_rel_x_u"] == 0.0 and out["max_rel_fields"] == 0.0
    # a perturbed result is flagged
    state["particles"]["ux"] = state["particles"]["ux"] * (1.0 + 1e-8)
    bad = bench.gpu_parity_leg(oracle, state, 200, 0)
    oracle_cdll.opal_ref_set_threads(1)      # cpu_reference_run raised the (process-wide) thread count of the shared library
    assert not bad["ok"] and bad["max_rel_x_u"] > 1e-10

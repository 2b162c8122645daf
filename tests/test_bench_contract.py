"""bench.py's CPU-only pieces: the reference arm prints one JSON line with the contract's keys
(bounded sample, all host cores), frame ids map to (replica, frame) as the weak-scaling run
needs, and the e2e byte counts follow the tensors that are copied."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_reference_arm_line():
    env = dict(os.environ, PCSEG_BENCH_REF_SAMPLE="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["unit"] == "Mpoints/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "Mpoints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    for k in ("metric", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "dtype", "config"):
        assert k in line, k


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_frame_ids_are_replicas_of_the_c2_sweep():
    import bench
    import synth
    a = bench._gen_frame(5)
    assert np.array_equal(a, synth.c2_frame(5, 256, bench.ROWS, bench.COLS), equal_nan=True)
    b = bench._gen_frame(256 + 5)   # rank 1's frame 5: same camera pose and boxes, other noise seeds
    assert b.shape == a.shape and not np.array_equal(a, b, equal_nan=True)
    assert abs(float(np.isnan(a).mean()) - float(np.isnan(b).mean())) < 0.01

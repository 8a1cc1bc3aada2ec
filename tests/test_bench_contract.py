"""bench.py's CPU arm (`--impl reference`) on a small sample: the JSON line carries the keys the driver reads and describes the same
workload, configuration keys and episode segment as the GPU arm.  CPU-only (the GPU arm itself runs on the GPU box)."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run_bench(*argv, env=None):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *argv], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, **(env or {})))
    assert p.returncode == 0, p.stderr[-2000:]
    return p.stdout


def test_reference_arm_line():
    out = run_bench("--impl", "reference", "--envs", "64", "--steps", "1", "--warmup", "3", "--substeps", "100")
    lines = [l for l in out.splitlines() if l.strip()]
    assert len(lines) == 1, out
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "rover env-steps/sec" and d["unit"] == "env-steps/s" and d["higher_is_better"] is True
    assert d["config"]["workload"].startswith("config3") and d["config"]["substeps_per_step"] == 100 and d["config"]["envs_per_gpu"] == 64
    assert d["steps"] == 1 and d["warmup"] == 3 and d["value"] > 0 and d["gpu_launches"] == 0 and d["scaling"] == "weak"
    assert d["e2e"] == {"value": d["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    c = d["cpu_baseline"]
    assert c["kind"] == "port" and c["cores"] == (os.cpu_count() or 1) and c["value"] == d["value"] and abs(c["per_core"] * c["cores"] - c["value"]) < 1e-6 * c["value"]
    assert "episode sub-steps 0..100 from reset" in c["sample"] and "64 envs" in c["sample"]


def test_both_arms_print_the_same_config_keys():
    """The driver compares the two arms' `config` objects key by key (same_config was false for a key mismatch)."""
    import bench
    a = bench.config_block("config3", 16384, 1, 500, 20)
    out = run_bench("--impl", "reference", "--steps", "1", "--substeps", "20")
    b = json.loads(out.strip())["config"]
    assert set(a) == set(b)
    assert b["envs_per_gpu"] == 16384 and b["workload"] == a["workload"]


def test_multi_gpu_default_is_the_config5_sweep_and_rank_0_alone_prints():
    out = run_bench("--impl", "reference", "--gpus", "2", "--steps", "1", "--substeps", "10", env={"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"})
    d = json.loads(out.strip())
    assert d["config"]["workload"].startswith("config5") and d["config"]["envs_per_gpu"] == (1 << 20) // 2 and d["config"]["envs_total"] == 1 << 20
    assert d["scaling"] == "strong" and d["n_gpus"] == 2
    out = run_bench("--impl", "reference", "--gpus", "2", "--envs", "64", "--steps", "1", env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert out.strip() == ""


def test_action_tape_is_a_function_of_step_and_global_id():
    import bench
    full = bench.action_tape(0, 3, 4096, 0)
    shard = bench.action_tape(1, 3, 1024, 2048)
    assert np.array_equal(full[2048:3072], shard)
    assert np.abs(full[:, 0]).max() <= 15 and np.abs(full[:, 1]).max() <= 25 and full[:, 0].std() > 5
    assert not np.array_equal(full, bench.action_tape(0, 4, 4096, 0))

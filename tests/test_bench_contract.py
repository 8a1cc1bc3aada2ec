"""bench.py's CPU arm (`--impl reference`) on a small sample: the JSON line carries the keys the driver reads and describes the same
episode segment as the GPU arm.  CPU-only (the GPU arm itself runs on the GPU box)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(*argv, env=None):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *argv], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, **(env or {})))
    assert p.returncode == 0, p.stderr[-2000:]
    return p.stdout


def test_reference_arm_line():
    out = run_bench("--impl", "reference", "--envs", "64", "--steps", "1", "--warmup", "3")
    lines = [l for l in out.splitlines() if l.strip()]
    assert len(lines) == 1, out
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "rover env-steps/sec" and d["unit"] == "env-steps/s" and d["higher_is_better"] is True
    assert d["config"]["workload"].startswith("config2") and d["config"]["substeps_per_step"] == 100 and d["config"]["envs"] == 64
    assert d["steps"] == 1 and d["warmup"] == 3 and d["value"] > 0 and d["gpu_launches"] == 0
    assert d["e2e"] == {"value": d["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    c = d["cpu_baseline"]
    assert c["kind"] == "port" and c["cores"] == (os.cpu_count() or 1) and c["value"] == d["value"]
    assert "episode sub-steps 300..400 from reset" in c["sample"] and "64 envs" in c["sample"]


def test_reference_arm_runs_on_rank_0_only():
    out = run_bench("--impl", "reference", "--gpus", "2", "--envs", "64", "--steps", "1", env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert out.strip() == ""

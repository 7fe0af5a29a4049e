"""bench.py's driver contract, as far as it can be checked without a GPU: the reference arm
runs on the host cores and prints ONE JSON line with the agreed keys; the GPU arm refuses to
run without a CUDA device instead of falling back."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args],
                          capture_output=True, text=True, timeout=600, env=e, cwd=ROOT)


def test_reference_arm_line():
    p = _run("--impl", "reference", "--steps", "1", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    assert d["metric"] == "aggregated+applied gradient GB/s" and d["unit"] == "GB/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["dtype"] == "f32" and d["data"] == "synthetic"
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0}
    assert d["value"] > 0


def test_reference_arm_other_ranks_are_silent():
    p = _run("--impl", "reference", "--gpus", "2", "--steps", "1",
             env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_gpu_arm_refuses_to_run_without_a_gpu():
    if os.path.exists("/dev/nvidiactl"):
        import pytest
        pytest.skip("box has a GPU")
    p = _run("--steps", "1", "--warmup", "1", "--no-e2e", "--no-cpu")
    assert p.returncode != 0
    assert not [ln for ln in p.stdout.splitlines() if ln.startswith("{")]


def test_reference_arm_runs_the_same_config_at_n_gpus():
    """At N > 1 the reference arm times the CPU path on the SAME N-PS-task workload the GPU arm
    runs (round 1's arm kept the single-PS job, so the driver's ratio compared unlike things)."""
    p = _run("--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1",
             env={"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"})
    assert p.returncode == 0, p.stderr[-2000:]
    d = json.loads([ln for ln in p.stdout.splitlines() if ln.startswith("{")][0])
    sys.path.insert(0, ROOT)
    import bench
    assert d["config"] == bench.workload_config(2)
    assert d["config"]["ps_shards"] == 2 and d["n_gpus"] == 2
    assert "2 PS tasks back to back" in d["cpu_baseline"]["sample"]
    assert d["cpu_baseline"]["host_mem_gbs"] > 0

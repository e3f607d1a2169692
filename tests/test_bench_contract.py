"""bench.py's one-line JSON contract, checked on the arm that runs without a GPU (CPU only)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*args, env=None):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, capture_output=True, text=True,
                          timeout=600, env={**os.environ, **(env or {})})


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    out = run("--impl", "reference", "--rows", "20000", "--steps", "2", "--warmup", "1", "--nq", "32")
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    j = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "impl", "cpu_baseline", "e2e"):
        assert key in j, key
    assert j["impl"] == "reference" and j["unit"] == "queries/s" and j["value"] > 0 and j["vs_baseline"] is None
    assert j["warmup"] >= 3  # the contract's floor is enforced whatever the caller asks for
    assert j["cpu_baseline"]["kind"] in ("port", "reference") and j["cpu_baseline"]["cores"] >= 1
    assert j["e2e"] == {"value": j["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in j["config"] and "model" not in j["config"]


def test_reference_arm_under_torchrun_only_rank0_prints():
    out = run("--impl", "reference", "--gpus", "2", "--rows", "5000", "--steps", "1", "--nq", "8", env={"RANK": "1", "WORLD_SIZE": "2"})
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_own_arm_fails_loudly_without_a_gpu():
    import torch

    if torch.cuda.is_available():
        return  # on a GPU box the -m gpu tests and the driver exercise this arm
    out = run("--steps", "1", "--no-cpu-baseline", "--no-scaling-base")
    assert out.returncode != 0
    assert "CUDA" in (out.stderr + out.stdout)

"""bench.py's CPU arm runs without a GPU: its JSON line must carry the keys the driver reads, and the GPU arm must
refuse to run without a device (no CPU fallback behind the product path)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=300, cwd=ROOT)


def test_reference_arm_line_has_the_contract_keys():
    p = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--width", "320", "--height", "200", "--ref-crop", "40", "20")
    assert p.returncode == 0, p.stderr
    line = json.loads(p.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Mrays/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["steps"] == 1 and line["n_gpus"] == 1
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1 and "crop" in line["cpu_baseline"]["sample"]
    assert line["e2e"] == {"value": line["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "spheres1024 320x200" in line["config"]["workload"]


def test_reference_arm_under_torchrun_env_runs_on_rank_0_only():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_gpu_arm_refuses_to_run_without_a_device():
    import torch
    if torch.cuda.is_available():
        return                      # on a GPU box the arm runs; the refusal is what the CPU box must show
    p = _run("--steps", "1", "--warmup", "0", "--width", "64", "--height", "40", "--no-cpu-baseline")
    assert p.returncode != 0
    assert "no CPU fallback" in (p.stderr + p.stdout)

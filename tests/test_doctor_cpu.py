"""python -m torch_distlearn_b200.doctor: the environment report and the CPU self-test."""
import json
import subprocess
import sys


def test_doctor_reports_and_selftests():
    out = subprocess.run([sys.executable, "-m", "torch_distlearn_b200.doctor", "--selftest", "--json"], capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    rep = json.loads(out.stdout.strip().splitlines()[-1])
    assert rep["libraries"]["host"]["built"] and rep["libraries"]["host"]["loads"]
    assert rep["selftest"]["host"]["ok"] is True
    assert rep["problems"] == []
    assert isinstance(rep["gpus"], list) and rep["cuda_available"] == bool(rep["gpus"])

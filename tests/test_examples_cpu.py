"""The example programs run end to end on CPU nodes (shared-memory transport), tiny synthetic data:
the launchers of the reference double as its smoke tests (SURVEY §4 "Manual tests")."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EX = os.path.join(ROOT, "examples")


def _spawn(script, n, port, *args, timeout=420):
    from torch_distlearn_b200.launch import spawn
    env = {"DISTLEARN_JOB": f"t{os.getpid()}", "PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    return spawn(os.path.join(EX, script), list(args), n, port=port, extra_env=env, timeout=timeout)


def test_mnist_allreduce_sgd_two_cpu_processes(tmp_path):
    """BASELINE.json config 1: MNIST convnet AllReduceSGD, 2 processes on CPU."""
    metrics = tmp_path / "m.jsonl"
    rc = _spawn("mnist.py", 2, 8201, "--numEpochs", "2", "--trainSamples", "600", "--batchSize", "4", "--save", str(tmp_path),
                "--metrics", str(metrics))
    assert rc == 0
    import json
    rows = [json.loads(l) for l in open(metrics)]
    assert len(rows) == 2 and rows[-1]["accuracy"] > rows[0]["accuracy"] and rows[-1]["accuracy"] > 0.8
    assert (tmp_path / "mnist.dlb200").exists()
    # resume from the checkpoint: starts at epoch 3
    rc = _spawn("mnist.py", 2, 8202, "--numEpochs", "3", "--trainSamples", "600", "--batchSize", "4", "--resume",
                str(tmp_path / "mnist.dlb200"), "--metrics", str(metrics), "--no-fused")
    assert rc == 0
    rows = [json.loads(l) for l in open(metrics)]
    assert len(rows) == 3 and rows[-1]["epoch"] == 3 and rows[-1]["accuracy"] > 0.9


def test_mnist_elastic_averaging_three_cpu_processes():
    assert _spawn("mnist_ea.py", 3, 8203, "--numEpochs", "1", "--trainSamples", "450", "--batchSize", "4", "--tau", "5") == 0


def test_async_easgd_server_clients_tester(tmp_path):
    env = dict(os.environ, EPOCHS="1", TRAIN="256", TEST="64", BATCH="32", TAU="2", TESTTIME="2", SAVE=str(tmp_path),
               DISTLEARN_JOB=f"t{os.getpid()}", PYTHONPATH=ROOT)
    res = subprocess.run(["bash", os.path.join(EX, "AsyncEASGD.sh"), "8204", "2"], env=env, capture_output=True, text=True,
                         timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "all clients finished after 4 syncs" in res.stdout
    log = open(tmp_path / "ErrorRate.log").read().strip().splitlines()
    assert log[0].startswith("Training Error") and len(log) >= 2
    assert (tmp_path / "easgd_center.dlb200").exists()


def test_cifar10_two_cpu_processes(tmp_path):
    """examples/cifar10.py (C8): global batch split over the nodes, train + test pass per epoch, confusion matrix all-reduce."""
    metrics = tmp_path / "c.jsonl"
    rc = _spawn("cifar10.py", 2, 8205, "--numEpochs", "1", "--trainSamples", "192", "--testSamples", "64", "--batchSize", "16",
                "--learningRate", "0.01", "--metrics", str(metrics))
    assert rc == 0
    import json
    row = json.loads(open(metrics).read().strip().splitlines()[-1])
    assert row["epoch"] == 1 and 0.0 <= row["test_accuracy"] <= 1.0 and row["samples_per_s"] > 0


def test_client_remote_over_torch_distributed():
    """examples/client_remote.py (C15): the AllReduceEA flow with rendezvous on --host/--port (torch.distributed backend)."""
    rc = _spawn("client_remote.py", 2, 8206, "--numEpochs", "1", "--trainSamples", "128", "--testSamples", "64", "--batchSize", "16",
                "--learningRate", "0.01", "--tau", "2")
    assert rc == 0

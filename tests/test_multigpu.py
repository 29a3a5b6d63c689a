"""Launches tests/run_multigpu.py under torchrun on every GPU of the box (needs >= 2 GPUs): real
cross-process symmetric heaps, every kernel variant vs torch, and the dist_checks scenarios."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(nproc, *extra, timeout=900, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(29511 + nproc),
           os.path.join(ROOT, "tests", "run_multigpu.py"), *extra]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT, env=env)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_all_gpus():
    n = torch.cuda.device_count()
    n = 8 if n >= 8 else 4 if n >= 4 else 2
    res = _torchrun(n, "--quick")
    assert res.returncode == 0 and "MULTIGPU OK" in res.stdout, res.stdout[-4000:] + res.stderr[-4000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_bucket_geometry_is_bit_exact_across_kernel_variants():
    """With a table that sends the whole vector through the two-shot kernel and its buckets through the one-shot
    kernel, the parameters must still agree bit for bit (the update's rounding is pinned with explicit fmaf)."""
    env = dict(os.environ, DL_B200_VARIANT_TABLE=os.path.join(ROOT, "tests", "variant_table_mixed.json"))
    res = _torchrun(2, "--quick", "--scenarios", "sgd_bucketed", env=env)
    assert res.returncode == 0 and "MULTIGPU OK" in res.stdout, res.stdout[-4000:] + res.stderr[-4000:]

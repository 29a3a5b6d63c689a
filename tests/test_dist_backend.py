"""The torch.distributed backend (several hosts / mixed CPU+GPU nodes) runs the same scenarios as the
shared-memory and NVLink transports: the algorithm classes must not care which transport is underneath."""
import os
import socket

import pytest

import dist_checks
from mp_util import run_ranks


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, job, port, scenario, kwargs):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from torch_distlearn_b200 import DistCommunicator
    comm = DistCommunicator(heap_mb=4)
    try:
        return dist_checks.SCENARIOS[scenario][0](comm, rank, world, **kwargs)
    finally:
        comm.close()
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.parametrize("scenario,world,kwargs", [
    ("sgd_reference_test", 3, {"seed": 21}),
    ("sgd_fused_uneven", 2, {}),
    ("ea_reference_test", 2, {"seed": 22, "use_flat": True}),
    ("tree_and_metrics", 2, {}),
    ("async_ea", 3, {"exclusive": True, "with_tester": False}),
    ("async_ea", 4, {"exclusive": False, "with_tester": True}),
])
def test_scenarios_over_torch_distributed(scenario, world, kwargs):
    results = run_ranks(_worker, world, _free_port(), scenario, kwargs, timeout=240.0)
    dist_checks.SCENARIOS[scenario][1](results, world, **kwargs)

"""The reference's two distributed tests (test/test_AllReduceSGD.lua, test/test_AllReduceEA.lua) and the
scenarios it never tested (fused step with uneven partitions, AsyncEA server/clients/tester, Tree
facade + confusion matrix), run as N spawned CPU processes over the shared-memory transport.
The same scenarios run on real GPUs in tests/test_multigpu.py."""
import pytest

import dist_checks
from mp_util import run_ranks


def _worker(rank, world, job, scenario, kwargs):
    import torch
    from torch_distlearn_b200 import HostCommunicator
    torch.manual_seed(0)
    comm = HostCommunicator(rank, world, name=job, heap_mb=8, timeout_s=60)
    try:
        return dist_checks.SCENARIOS[scenario][0](comm, rank, world, **kwargs)
    finally:
        comm.close()


def _run(scenario, world, **kwargs):
    results = run_ranks(_worker, world, scenario, kwargs, timeout=240.0)
    dist_checks.SCENARIOS[scenario][1](results, world, **kwargs)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_allreduce_sgd_reference_property(world):
    """numNodes in {2,4,8}, random uneven step counts, exact equality (test/test_AllReduceSGD.lua:24-39)."""
    _run("sgd_reference_test", world, seed=world)


@pytest.mark.parametrize("world", [2, 3])
def test_fused_sgd_uneven_steps(world):
    _run("sgd_fused_uneven", world)


@pytest.mark.parametrize("world,use_flat", [(2, False), (4, True), (8, False)])
def test_allreduce_ea_reference_property(world, use_flat):
    """tau=3, alpha=0.4, random walk with halving steps, nodes end < 1e-6 apart (test/test_AllReduceEA.lua)."""
    _run("ea_reference_test", world, seed=10 + world, use_flat=use_flat)


@pytest.mark.parametrize("world,exclusive,with_tester", [(3, True, False), (4, False, True), (4, True, True)])
def test_async_ea_roles(world, exclusive, with_tester):
    _run("async_ea", world, exclusive=exclusive, with_tester=with_tester)


def test_tree_facade_and_confusion_matrix():
    _run("tree_and_metrics", 3)


@pytest.mark.parametrize("world", [2, 3])
def test_bucketed_step_equals_whole_vector_step(world):
    """Per-bucket launches from autograd hooks (uneven steps, drain rounds) give bit-identical parameters."""
    _run("sgd_bucketed", world)


@pytest.mark.parametrize("world", [2, 3])
def test_table_api_with_an_empty_partition(world):
    """A node that took no step reaches synchronizeParameters before it ever staged a gradient table: the staging
    buffers must still sit at the same heap offsets on every node (lua/AllReduceSGD.lua:33-54)."""
    _run("sgd_table_empty_partition", world)

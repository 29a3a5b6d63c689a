"""Multi-rank scenarios shared by the CPU tests (HostCommunicator, spawned processes) and the
multi-GPU tests (P2PCommunicator under torchrun).  Each scenario has a ``run_*`` half that executes
on every rank and returns CPU tensors, and a ``verify_*`` half that checks all ranks' results
against the oracle.  They port the reference's two tests (test/test_AllReduceSGD.lua,
test/test_AllReduceEA.lua) and add what the reference never tested (AsyncEA, fused paths)."""
from __future__ import annotations

import random
from typing import Any, Dict, List

import torch

import oracle


def _dev(comm):
    return comm.device


def _steps_for(rank: int, epoch: int, lo: int, hi: int, seed: int) -> int:
    return random.Random(seed * 1000003 + epoch * 1009 + rank).randint(lo, hi)


# ---- AllReduceSGD: the reference unit test (test/test_AllReduceSGD.lua:8-21) --------------------------
def run_sgd_reference_test(comm, rank, world, seed=0, epochs=3, lo=4, hi=13) -> Dict[str, Any]:
    import torch_distlearn_b200 as dl
    sgd = dl.AllReduceSGD(comm)
    dev = _dev(comm)
    params = [torch.randn(7, generator=torch.Generator().manual_seed(seed + rank)).to(dev)]
    grads = [torch.zeros(7, device=dev)]
    sgd.synchronizeParameters(params)
    for epoch in range(epochs):
        steps = _steps_for(rank, epoch, lo, hi, seed)
        for _ in range(steps):
            grads[0].fill_(1.0 / steps)
            sgd.sumAndNormalizeGradients(grads)
            params[0].add_(grads[0])
        sgd.synchronizeParameters(params)
    return {"params": params[0].cpu()}


def verify_sgd_reference_test(results: List[Dict[str, Any]], world, seed=0, epochs=3, lo=4, hi=13) -> None:
    for r in range(1, world):
        assert torch.equal(results[0]["params"], results[r]["params"]), "all nodes should have identical params"
    sim = oracle.SgdNodeSim([torch.randn(7, generator=torch.Generator().manual_seed(seed + r)) for r in range(world)])
    sim.synchronize([0] * world)
    for epoch in range(epochs):
        steps = [_steps_for(r, epoch, lo, hi, seed) for r in range(world)]
        sim.epoch(steps, lambda r, s, k: torch.full((7,), 1.0 / k))
    torch.testing.assert_close(results[0]["params"], sim.params[0], rtol=1e-5, atol=1e-5)


# ---- AllReduceSGD fused step with uneven steps -----------------------------------------------------------
def run_sgd_fused_uneven(comm, rank, world, seed=1, epochs=2, n=1000) -> Dict[str, Any]:
    import torch_distlearn_b200 as dl
    dev = _dev(comm)
    p0 = torch.randn(n, generator=torch.Generator().manual_seed(seed + 17 * rank)).to(dev)
    flat = dl.FlatParams(comm, [p0.clone().requires_grad_(True)], with_momentum=True)
    sgd = dl.AllReduceSGD(comm, flat)
    sgd.synchronizeParameters()
    for epoch in range(epochs):
        steps = _steps_for(rank, epoch, 2, 6, seed)
        for s in range(steps):
            g = torch.randn(n, generator=torch.Generator().manual_seed(seed + 7919 * epoch + 101 * s + rank)).to(dev)
            flat.grads.tensor[:n].copy_(g)
            sgd.step(lr=0.1 / (1 + epoch), momentum=0.5)
        winner = sgd.synchronizeParameters()
    comm.synchronize()
    return {"params": flat.params.tensor[:n].cpu().clone(), "grads_max": float(flat.grads.tensor.abs().max()), "winner": winner}


def verify_sgd_fused_uneven(results, world, seed=1, epochs=2, n=1000) -> None:
    p = torch.randn(n, generator=torch.Generator().manual_seed(seed))      # rank 0's init wins the first sync
    m = torch.zeros(n)
    last_steps = None
    for epoch in range(epochs):
        steps = [_steps_for(r, epoch, 2, 6, seed) for r in range(world)]
        last_steps = steps
        for s in range(max(steps)):
            active = [1 if s < k else 0 for k in steps]
            grads = [torch.randn(n, generator=torch.Generator().manual_seed(seed + 7919 * epoch + 101 * s + r)) for r in range(world)]
            p, m, _ = oracle.fused_sgd(grads, p, 0.1 / (1 + epoch), active, 0.5, m)
    for r in range(world):
        assert torch.equal(results[r]["params"], results[0]["params"]), "bit-identical parameters across ranks"
        assert results[r]["grads_max"] == 0.0
        assert results[r]["winner"] == oracle.winner(last_steps)
    torch.testing.assert_close(results[0]["params"], p, rtol=1e-4, atol=1e-5)


# ---- AllReduceEA: the reference unit test (test/test_AllReduceEA.lua:7-20) ------------------------------
def run_ea_reference_test(comm, rank, world, seed=2, epochs=3, lo=45, hi=53, tau=3, alpha=0.4, use_flat=False):
    import torch_distlearn_b200 as dl
    dev = _dev(comm)
    x0 = torch.randn(7, generator=torch.Generator().manual_seed(seed + rank))
    if use_flat:
        flat = dl.FlatParams(comm, [x0.clone().to(dev).requires_grad_(True)], with_grads=False)
        ea = dl.AllReduceEA(comm, tau, alpha, flat=flat)
        params = flat
        x = flat.params.tensor[:7]
    else:
        ea = dl.AllReduceEA(comm, tau, alpha)
        params = [x0.clone().to(dev)]
        x = params[0]
    ea.synchronizeParameters(params)
    gen = torch.Generator().manual_seed(seed * 31 + rank)
    slowit = 1.0
    for epoch in range(epochs):
        steps = _steps_for(rank, epoch, lo, hi, seed)
        for _ in range(steps):
            x.add_((torch.randn(7, generator=gen) / slowit).to(dev))     # wander
            ea.averageParameters(params)
            slowit *= 2
        ea.synchronizeCenter(params)
    comm.synchronize()
    return {"params": x.cpu().clone(), "center": ea.center[:7].cpu().clone()}


def verify_ea_reference_test(results, world, seed=2, epochs=3, lo=45, hi=53, tau=3, alpha=0.4, **_kw) -> None:
    for r in range(1, world):
        delta = float((results[0]["params"] - results[r]["params"]).abs().max())
        assert delta < 1e-6, f"all nodes should be really close together {delta}"
        assert torch.equal(results[0]["center"], results[r]["center"]), "center replicas are bit-identical here"
    # full oracle replay
    xs = [torch.randn(7, generator=torch.Generator().manual_seed(seed + r)) for r in range(world)]
    sim = oracle.EaNodeSim(xs, tau, alpha)
    sim.synchronize_parameters()
    gens = [torch.Generator().manual_seed(seed * 31 + r) for r in range(world)]
    slow = [1.0] * world

    def move(r, _i):
        v = torch.randn(7, generator=gens[r]) / slow[r]
        slow[r] *= 2
        return v

    for epoch in range(epochs):
        steps = [_steps_for(r, epoch, lo, hi, seed) for r in range(world)]
        sim.epoch(steps, move)
    for r in range(world):
        torch.testing.assert_close(results[r]["params"], sim.xs[r], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(results[0]["center"], sim.center, rtol=1e-4, atol=1e-5)


# ---- AsyncEA: server + clients (+ tester) ---------------------------------------------------------------
def run_async_ea(comm, rank, world, exclusive=True, tau=2, alpha=0.25, local_steps=6, n=37, with_tester=False):
    import torch_distlearn_b200 as dl
    dev = _dev(comm)
    tester_rank = world - 1 if with_tester else None
    num_clients = world - 1 - (1 if with_tester else 0)
    role = "server" if rank == 0 else ("tester" if rank == tester_rank else "client")
    algo = dl.AsyncEA(comm, role, tau, alpha, server_rank=0, num_clients=num_clients, exclusive=exclusive)
    params = {"w": torch.full((n,), float(rank), device=dev), "b": [torch.full((3,), 10.0 + rank, device=dev)]}
    out: Dict[str, Any] = {"role": role}
    if role == "server":
        params["w"].fill_(100.0)
        algo.initServer(params)
        served = 0
        tests = 0
        while algo.syncServer(params, timeout=60.0):
            served += 1
            if with_tester and tests == 0 and algo.syncs >= 1:
                algo.testNet(block=True, timeout=60.0)
                tests += 1
        algo.stop()
        out.update(center_w=params["w"].cpu().clone(), center_b=params["b"][0].cpu().clone(), syncs=algo.syncs, tests=tests)
    elif role == "client":
        algo.initClient(params)
        assert float(params["w"][0]) == 100.0, "clients start from the server's center"
        synced = 0
        for s in range(local_steps):
            params["w"].add_(1.0)                      # "training"
            synced += bool(algo.syncClient(params))
        algo.finish()
        out.update(w=params["w"].cpu().clone(), b=params["b"][0].cpu().clone(), synced=synced)
    else:
        algo.initTester(params)
        seen = 0
        while algo.startTest(params, timeout=60.0):
            seen += 1
            out["tested_w"] = params["w"].cpu().clone()
            algo.finishTest()
        out["tests_seen"] = seen
    comm.barrier()
    return out


def verify_async_ea(results, world, exclusive=True, tau=2, alpha=0.25, local_steps=6, n=37, with_tester=False) -> None:
    server = results[0]
    clients = [r for r in results if r["role"] == "client"]
    assert server["syncs"] == len(clients) * (local_steps // tau)
    for c in clients:
        assert c["synced"] == local_steps // tau
    # conservation: every sync moves d out of a client and into the center
    total_w = server["center_w"].double() + sum(c["w"].double() for c in clients)
    want = 100.0 + len(clients) * (100.0 + local_steps)
    torch.testing.assert_close(total_w, torch.full((n,), want, dtype=torch.float64), rtol=1e-5, atol=1e-4)
    assert float(server["center_w"].max() - server["center_w"].min()) < 1e-4
    if with_tester:
        tester = [r for r in results if r["role"] == "tester"][0]
        assert tester["tests_seen"] == server["tests"] == 1
        assert tester["tested_w"].shape[0] == n


# ---- Tree facade + confusion matrix -------------------------------------------------------------------------
def run_tree_and_metrics(comm, rank, world):
    import torch_distlearn_b200 as dl
    from torch_distlearn_b200.utils import ConfusionMatrix
    dev = _dev(comm)
    tree = dl.Tree(comm)
    assert tree.nodeIndex == rank + 1 and tree.numNodes == world
    mat = torch.full((10, 10), rank + 1, dtype=torch.int64, device=dev)
    _, n = tree.allReduce(mat, lambda a, b: a.add_(b))
    table = {"a": torch.full((5,), float(rank), device=dev), "b": [torch.full((2, 2), float(rank) * 2, device=dev)]}
    tree.scatter(table)
    cm = ConfusionMatrix([str(i) for i in range(10)], comm=comm)
    logits = torch.zeros(4, 10, device=dev)
    logits[:, rank % 10] = 1.0
    cm.batchAdd(logits, torch.full((4,), rank % 10, device=dev))
    cm.allReduce()
    comm.synchronize()
    return {"mat": mat.cpu(), "n": n, "a": table["a"].cpu(), "b": table["b"][0].cpu(), "cm": cm.mat.cpu().clone()}


def verify_tree_and_metrics(results, world) -> None:
    want = sum(r + 1 for r in range(world))
    for res in results:
        assert res["n"] == world
        assert torch.all(res["mat"] == want)
        assert torch.all(res["a"] == 0.0) and torch.all(res["b"] == 0.0)     # node 1's values
        assert int(res["cm"].sum()) == 4 * world and int(res["cm"].diag().sum()) == 4 * world



# ---- bucketed fused step (overlap with backward) == whole-vector fused step ----------------------------------
def _mlp(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(20, 64), torch.nn.Tanh(), torch.nn.Linear(64, 33), torch.nn.Tanh(),
                               torch.nn.Linear(33, 10), torch.nn.LogSoftmax(dim=1))


def run_sgd_bucketed(comm, rank, world, seed=5, epochs=2, use_graph=False) -> Dict[str, Any]:
    import torch.nn.functional as F
    import torch_distlearn_b200 as dl
    from torch_distlearn_b200.parallel.trainer import TrainStep
    dev = _dev(comm)
    outs = {}
    for mode, bucket_mb in (("whole", 0.0), ("bucketed", 0.001)):          # 0.001 MB -> ~260-element buckets
        model = _mlp(seed).to(dev)
        flat = dl.FlatParams(comm, model, with_momentum=True, name=mode)
        sgd = dl.AllReduceSGD(comm, flat, channel=0)
        sgd.synchronizeParameters()
        step = TrainStep(model, flat, sgd, (8, 20), lr=0.05, momentum=0.9, weight_decay=1e-3, use_graph=use_graph,
                         warmup=1, bucket_mb=bucket_mb, bucket_blocks=4)
        if bucket_mb:
            assert len(step.buckets.buckets) >= 3
        for epoch in range(epochs):
            steps = _steps_for(rank, epoch, 2, 5, seed) if not use_graph else 4
            for s in range(steps):
                g = torch.Generator().manual_seed(seed + 977 * epoch + 31 * s + rank)
                x = torch.randn(8, 20, generator=g)
                y = torch.randint(0, 10, (8,), generator=g)
                step.step(x.to(dev), y.to(dev))
            sgd.synchronizeParameters()
        comm.synchronize()
        outs[mode] = flat.params.tensor[:flat.numel].cpu().clone()
        outs[mode + "_grads_max"] = float(flat.grads.tensor.abs().max())
        if step.buckets is not None:
            step.buckets.remove()
    return outs


def verify_sgd_bucketed(results, world, **_kw) -> None:
    for r in range(world):
        assert torch.equal(results[r]["whole"], results[0]["whole"])
        assert torch.equal(results[r]["bucketed"], results[0]["bucketed"])
        assert results[r]["whole_grads_max"] == 0.0 and results[r]["bucketed_grads_max"] == 0.0
    # owner-computes with rank-order sums: the bucket geometry must not change a single bit
    assert torch.equal(results[0]["whole"], results[0]["bucketed"])
    assert torch.isfinite(results[0]["whole"]).all()



# ---- table API, one node with an empty partition (no initial synchronizeParameters) ---------------------------
def run_sgd_table_empty_partition(comm, rank, world, steps0=2) -> Dict[str, Any]:
    """Rank `world-1` takes no step at all and goes straight to synchronizeParameters with a plain table; the others
    stage gradients first.  Staging buffers must land at the same heap offsets on every rank regardless."""
    import torch_distlearn_b200 as dl
    sgd = dl.AllReduceSGD(comm)
    dev = _dev(comm)
    params = [torch.full((5,), 2.0 + rank).to(dev), torch.full((3, 2), -1.0 - rank).to(dev)]
    grads = [torch.zeros(5, device=dev), torch.zeros(3, 2, device=dev)]
    steps = 0 if rank == world - 1 else steps0 + rank
    for s in range(steps):
        for g in grads:
            g.fill_(0.1 * (s + 1))
        sgd.sumAndNormalizeGradients(grads)
        for p, g in zip(params, grads):
            p.sub_(g)
    winner = sgd.synchronizeParameters(params)
    return {"params": [p.cpu() for p in params], "winner": winner}


def verify_sgd_table_empty_partition(results, world, steps0=2) -> None:
    steps = [0 if r == world - 1 else steps0 + r for r in range(world)]
    best = max(steps)
    winner = max(r for r, s in enumerate(steps) if s == best)
    # the winner's own trajectory: at step s the mean over the ranks still active of 0.1 (s+1) is 0.1 (s+1)
    want = [torch.full((5,), 2.0 + winner), torch.full((3, 2), -1.0 - winner)]
    for s in range(best):
        for w in want:
            w.sub_(0.1 * (s + 1))
    for r in range(world):
        assert results[r]["winner"] == winner
        for got, w in zip(results[r]["params"], want):
            torch.testing.assert_close(got, w, rtol=1e-6, atol=1e-6)
        for a, b in zip(results[r]["params"], results[0]["params"]):
            assert torch.equal(a, b), "every node must end with the winner's bits"

SCENARIOS = {
    "sgd_reference_test": (run_sgd_reference_test, verify_sgd_reference_test),
    "sgd_fused_uneven": (run_sgd_fused_uneven, verify_sgd_fused_uneven),
    "ea_reference_test": (run_ea_reference_test, verify_ea_reference_test),
    "async_ea": (run_async_ea, verify_async_ea),
    "tree_and_metrics": (run_tree_and_metrics, verify_tree_and_metrics),
    "sgd_bucketed": (run_sgd_bucketed, verify_sgd_bucketed),
    "sgd_table_empty_partition": (run_sgd_table_empty_partition, verify_sgd_table_empty_partition),
}

"""CPU shared-memory transport vs the oracle (multi-process, no GPU).  The host transport runs the
same algorithms as the CUDA kernels, so this also pins the cross-rank protocol (flags, payloads,
active/inactive ranks, owner-computes slices)."""
import pytest
import torch

import oracle
from mp_util import run_ranks


def _mk(rank, world, job, heap_mb=8):
    from torch_distlearn_b200 import HostCommunicator
    return HostCommunicator(rank, world, name=job, heap_mb=heap_mb, timeout_s=30)


def _w_allreduce(rank, world, job, n, active, normalize, dtype_name):
    dtype = getattr(torch, dtype_name)
    comm = _mk(rank, world, job)
    buf = comm.alloc(n, dtype)
    g = torch.Generator().manual_seed(100 + rank)
    if dtype.is_floating_point:
        buf.tensor.copy_(torch.randn(n, generator=g, dtype=torch.float64).to(dtype))
    else:
        buf.tensor.copy_(torch.randint(-1000, 1000, (n,), generator=g).to(dtype))
    res = comm.allreduce(buf, n, active=active[rank], payload=7 + rank, normalize=normalize)
    out = buf.tensor.clone()
    comm.close()
    return out, res.n_active, res.payloads


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("dtype_name", ["float32", "int64", "int32", "float64"])
def test_allreduce_matches_oracle(world, dtype_name):
    n = 1000 if dtype_name in ("float32", "int32") else 502
    active = [1] * world
    outs = run_ranks(_w_allreduce, world, n, active, False, dtype_name)
    dtype = getattr(torch, dtype_name)
    vals = []
    for r in range(world):
        g = torch.Generator().manual_seed(100 + r)
        vals.append(torch.randn(n, generator=g, dtype=torch.float64).to(dtype) if dtype.is_floating_point
                    else torch.randint(-1000, 1000, (n,), generator=g).to(dtype))
    want, n_act = oracle.allreduce_sum(vals)
    for r in range(world):
        out, na, pay = outs[r]
        assert na == n_act == world
        assert pay == [7 + k for k in range(world)]
        assert torch.equal(out, outs[0][0]), "all ranks must hold identical bits"
        if dtype.is_floating_point:
            torch.testing.assert_close(out, want, rtol=1e-6, atol=1e-6)
        else:
            assert torch.equal(out, want)


def test_allreduce_inactive_ranks_contribute_zero_and_normalise():
    world, n = 4, 64
    active = [1, 0, 1, 0]
    outs = run_ranks(_w_allreduce, world, n, active, True, "float32")
    vals = [torch.randn(n, generator=torch.Generator().manual_seed(100 + r), dtype=torch.float64).float() for r in range(world)]
    want, n_act = oracle.allreduce_sum(vals, active, normalize=True)
    assert n_act == 2
    for out, na, _ in outs:
        assert na == 2
        torch.testing.assert_close(out, want, rtol=1e-6, atol=1e-6)


def _w_fused_sgd(rank, world, job, n, active, lr, momentum, wd, nesterov, steps):
    comm = _mk(rank, world, job)
    g_buf, p_buf = comm.alloc(n), comm.alloc(n)
    m_buf = comm.alloc(n) if momentum else None
    p_buf.tensor.copy_(torch.randn(n, generator=torch.Generator().manual_seed(5)))
    n_acts = []
    for s in range(steps):
        g_buf.tensor.copy_(torch.randn(n, generator=torch.Generator().manual_seed(1000 * s + rank)))
        res = comm.fused_sgd(g_buf, p_buf, m_buf, n, lr, momentum=momentum, weight_decay=wd, nesterov=nesterov,
                             active=active[s][rank], payload=s)
        n_acts.append(res.n_active)
        assert float(g_buf.tensor.abs().max()) == 0.0, "gradients must be zeroed by the fused step"
    out = p_buf.tensor.clone()
    comm.close()
    return out, n_acts


@pytest.mark.parametrize("world,momentum,wd,nesterov", [(2, 0.0, 0.0, False), (4, 0.9, 1e-3, False), (3, 0.9, 0.0, True)])
def test_fused_sgd_matches_oracle(world, momentum, wd, nesterov):
    n, lr, steps = 256, 0.05, 4
    active = [[1] * world, [1] + [0] * (world - 1), [0] * world, [1] * world]
    outs = run_ranks(_w_fused_sgd, world, n, active, lr, momentum, wd, nesterov, steps)
    p = torch.randn(n, generator=torch.Generator().manual_seed(5))
    m = torch.zeros(n) if momentum else None
    for s in range(steps):
        grads = [torch.randn(n, generator=torch.Generator().manual_seed(1000 * s + r)) for r in range(world)]
        p, m, n_act = oracle.fused_sgd(grads, p, lr, active[s], momentum, m, wd, nesterov)
        for out, n_acts in outs:
            assert n_acts[s] == n_act
    for out, _ in outs:
        assert torch.equal(out, outs[0][0]), "parameters must be bit-identical across ranks"
        torch.testing.assert_close(out, p, rtol=1e-5, atol=1e-6)


def _w_fused_ea(rank, world, job, n, alpha, rounds):
    comm = _mk(rank, world, job)
    x, c0, c1 = comm.alloc(n), comm.alloc(n), comm.alloc(n)
    x.tensor.copy_(torch.randn(n, generator=torch.Generator().manual_seed(10 + rank)))
    c0.tensor.copy_(torch.randn(n, generator=torch.Generator().manual_seed(3)))
    cur = [c0, c1]
    for r in range(rounds):
        comm.fused_ea(x, cur[0], cur[1], n, alpha, active=1)
        cur.reverse()
        x.tensor.add_(0.01 * (rank + 1))
    out = (x.tensor.clone(), cur[0].tensor.clone())
    comm.close()
    return out


@pytest.mark.parametrize("world", [2, 4])
def test_fused_ea_matches_oracle(world):
    n, alpha, rounds = 128, 0.2, 3
    outs = run_ranks(_w_fused_ea, world, n, alpha, rounds)
    xs = [torch.randn(n, generator=torch.Generator().manual_seed(10 + r)) for r in range(world)]
    c = torch.randn(n, generator=torch.Generator().manual_seed(3))
    for _ in range(rounds):
        xs, c = oracle.elastic_average(xs, c, alpha)
        xs = [x + 0.01 * (r + 1) for r, x in enumerate(xs)]
    for r, (x, cc) in enumerate(outs):
        torch.testing.assert_close(x, xs[r], rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(cc, c, rtol=1e-5, atol=1e-6)
        assert torch.equal(cc, outs[0][1]), "center replicas must be bit-identical"


def _w_broadcast(rank, world, job, n, root, with_also):
    comm = _mk(rank, world, job)
    a, b = comm.alloc(n), comm.alloc(n)
    a.tensor.fill_(float(rank + 1))
    b.tensor.fill_(-1.0)
    res = comm.broadcast(a, n, root=root, also=b if with_also else None, payload=rank)
    out = (a.tensor.clone(), b.tensor.clone(), res.payloads)
    comm.close()
    return out


@pytest.mark.parametrize("world,root,with_also", [(2, 1, False), (4, 2, True), (3, 0, True)])
def test_broadcast(world, root, with_also):
    n = 100
    outs = run_ranks(_w_broadcast, world, n, root, with_also)
    for a, b, pay in outs:
        assert torch.all(a == float(root + 1))
        assert torch.all(b == (float(root + 1) if with_also else -1.0))
        assert pay == list(range(world))


def _w_async(rank, world, job, n, alpha, exclusive, syncs):
    comm = _mk(rank, world, job)
    center = comm.alloc(n)
    if rank == 0:
        center.tensor.fill_(1.0)
        comm.mailbox_write(0, 0, 1)
        while comm.sync_count(0) < syncs * (world - 1):
            pass
        out = center.tensor.clone()
    else:
        while comm.mailbox_read(0, 0) != 1:
            pass
        x = torch.full((n,), float(rank + 1))
        for _ in range(syncs):
            comm.async_client_sync(0, x, center, n, alpha, exclusive)
        out = x.clone()
    comm.barrier()
    comm.close()
    return out


@pytest.mark.parametrize("exclusive", [True, False])
def test_async_ea_conserves_mass(exclusive):
    """Every sync moves d out of x and into the center: sum(x_i) + center is invariant, whatever
    the interleaving (exactly in exclusive mode, up to fp32 rounding in lock-free mode)."""
    world, n, alpha, syncs = 4, 64, 0.2, 5
    outs = run_ranks(_w_async, world, n, alpha, exclusive, syncs)
    total = sum(o.double() for o in outs)
    want = 1.0 + sum(float(r + 1) for r in range(1, world))
    torch.testing.assert_close(total, torch.full((n,), want, dtype=torch.float64), rtol=1e-5, atol=1e-5)
    # with the mutex every sync sees a whole center, so all elements are treated alike; lock-free
    # syncs may read a center that another client is half-way through updating (torn reads are
    # part of the contract), so only conservation holds there
    if exclusive:
        for o in outs:
            assert float(o.max() - o.min()) < 1e-5


def _w_timeout(rank, world, job):
    comm = _mk(rank, world, job)
    comm.set_timeout(0.5)
    if rank == 1:
        comm.close()
        return "skipped"
    try:
        comm.barrier()
    except Exception as e:  # noqa: BLE001
        comm.close()
        return type(e).__name__
    comm.close()
    return "no error"


def test_missing_rank_times_out_cleanly():
    """Fault injection: a rank that never launches must produce an error, not a hang."""
    outs = run_ranks(_w_timeout, 2)
    assert outs[0] == "NativeError" and outs[1] == "skipped"

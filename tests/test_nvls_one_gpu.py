"""The switch-reduction kernels on a ONE-GPU box: a world-1 communicator whose heap is VMM memory bound to a multicast
object of a single device (`P2PCommunicator(self_multicast=True)`), so `multimem.ld_reduce` / `multimem.st` — the nvls,
hybrid and light variants — execute where the driver's test machine can see them.  With one rank the sum is the rank's own
data, so every result has a closed form; what is tested is the addressing through the multicast view, the slicing and
the epilogues.  Multi-rank values are checked by tests/run_multigpu.py on 2/4/8 GPUs.

If the driver refuses a one-device multicast object the tests skip with its message (recorded in DESIGN.md §7)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def comm():
    from torch_distlearn_b200.parallel.comm import P2PCommunicator
    try:
        c = P2PCommunicator(rank=0, world=1, device=torch.device("cuda", 0), heap_mb=96, self_multicast=True)
    except RuntimeError as e:
        pytest.skip(f"no single-device multicast on this box: {e}")
    yield c
    c.close()


@pytest.fixture(autouse=True)
def heap_mark(comm):
    """Every test gives its buffers back (the heap has a bump allocator)."""
    mark = comm._cursor
    yield
    torch.cuda.synchronize()
    comm._cursor = mark


def test_the_heap_is_bound_to_a_multicast_object(comm):
    assert comm.has_multicast and comm.mode == "vmm"
    assert comm._c.mc, "multicast view is mapped"


@pytest.mark.parametrize("variant,permille", [("nvls", None), ("hybrid", 1000), ("hybrid", 500), ("light", 1000)])
@pytest.mark.parametrize("n", [4096, 1 << 20, (1 << 20) + 8 * 37])
def test_fused_sgd_through_the_switch(comm, variant, permille, n):
    g = comm.alloc(n, torch.float32, "g")
    p = comm.alloc(n, torch.float32, "p")
    m = comm.alloc(n, torch.float32, "m")
    gen = torch.Generator(device="cuda").manual_seed(n)
    g.tensor.copy_(torch.randn(n, device="cuda", generator=gen))
    p.tensor.copy_(torch.randn(n, device="cuda", generator=gen))
    m.tensor.copy_(torch.randn(n, device="cuda", generator=gen))
    g0, p0, m0 = g.tensor.double().clone(), p.tensor.double().clone(), m.tensor.double().clone()
    if permille is not None:
        comm._variants.hybrid_permille = {op: permille for op in comm._variants.OPS}
        comm._variants.light_permille = permille
    res = comm.fused_sgd(g, p, m, n, lr=0.05, momentum=0.9, weight_decay=1e-3, nesterov=False, variant=variant, read=True)
    torch.cuda.synchronize()
    assert res.n_active == 1
    gg = g0 + 1e-3 * p0
    mm = 0.9 * m0 + gg
    torch.testing.assert_close(m.tensor.double(), mm, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(p.tensor.double(), p0 - 0.05 * mm, rtol=1e-6, atol=1e-6)
    assert float(g.tensor.abs().max()) == 0.0, "gradients are zeroed by the same launch"


@pytest.mark.parametrize("variant,permille", [("nvls", None), ("hybrid", 1000), ("hybrid", 250)])
def test_allreduce_through_the_switch(comm, variant, permille):
    n = (1 << 19) + 64
    b = comm.alloc(n, torch.float32, "b")
    b.tensor.copy_(torch.arange(n, device="cuda", dtype=torch.float32) * 0.25 - 7.0)
    want = b.tensor.clone()
    if permille is not None:
        comm._variants.hybrid_permille = {op: permille for op in comm._variants.OPS}
    comm.allreduce(b, n, normalize=True, variant=variant)
    torch.cuda.synchronize()
    assert torch.equal(b.tensor, want), "one rank: the normalised sum is the input, bit for bit"


@pytest.mark.parametrize("variant,permille", [("nvls", None), ("hybrid", 1000)])
def test_fused_ea_through_the_switch(comm, variant, permille):
    n = 1 << 18
    x = comm.alloc(n, torch.float32, "x")
    c0 = comm.alloc(n, torch.float32, "c0")
    c1 = comm.alloc(n, torch.float32, "c1")
    gen = torch.Generator(device="cuda").manual_seed(5)
    x.tensor.copy_(torch.randn(n, device="cuda", generator=gen))
    c0.tensor.copy_(torch.randn(n, device="cuda", generator=gen))
    xs, cs = x.tensor.double().clone(), c0.tensor.double().clone()
    if permille is not None:
        comm._variants.hybrid_permille = {op: permille for op in comm._variants.OPS}
    comm.fused_ea(x, c0, c1, n, alpha=0.2, variant=variant)
    torch.cuda.synchronize()
    d = 0.2 * (xs - cs)
    torch.testing.assert_close(c1.tensor.double(), cs + d, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(x.tensor.double(), xs - d, rtol=1e-6, atol=1e-6)


def test_broadcast_through_the_switch(comm):
    n = 1 << 18
    a = comm.alloc(n, torch.float32, "a")
    also = comm.alloc(n, torch.float32, "also")
    a.tensor.copy_(torch.arange(n, device="cuda", dtype=torch.float32))
    comm.broadcast(a, n, root=0, also=also, variant="nvls")
    torch.cuda.synchronize()
    assert torch.equal(also.tensor, a.tensor)
    assert torch.equal(a.tensor, torch.arange(n, device="cuda", dtype=torch.float32))

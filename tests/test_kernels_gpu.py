"""sm_100a kernels vs a plain PyTorch fp32 reference, on ONE GPU.

Multi-rank kernels are exercised through EmulatedCommunicator: every emulated rank has its own
heap on the same device and one cooperative launch (gridDim.y = world) plays all ranks, so the
flag protocol, slicing, active masks and payloads run exactly as across GPUs (B200_PROFILING.md:
never run inter-dependent ranks as separate launches on one GPU)."""
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _emu(world, heap_mb=24):
    from torch_distlearn_b200 import EmulatedCommunicator
    return EmulatedCommunicator(world, torch.device(DEV), heap_mb=heap_mb)


def _rand(n, seed, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    if dtype.is_floating_point:
        return torch.randn(n, generator=g, dtype=torch.float64).to(dtype)
    return torch.randint(-1000, 1000, (n,), generator=g).to(dtype)


def test_library_is_native_and_loaded():
    from torch_distlearn_b200 import _lib
    assert _lib.cuda().dl_version() >= 100
    maps = open("/proc/self/maps").read()
    assert "libdistlearn_b200.so" in maps


@pytest.mark.parametrize("world", [2, 4, 8, 3])
@pytest.mark.parametrize("variant", ["oneshot", "twoshot", "hybrid", "ll"])
@pytest.mark.parametrize("n", [64, 4096 + 32, 300000])
def test_allreduce_f32(world, variant, n):
    if variant == "oneshot":
        n = min(n, 100000)   # one-shot keeps the whole vector in registers: capacity-bound by design
    comm = _emu(world)
    if variant == "ll":
        assert comm.ll_max_bytes >= 4096
        n = min(n, comm.ll_max_bytes // 4)   # the push path is for small messages: capped by its staging area
    buf, views = comm.alloc_all(n)
    vals = [_rand(n, 10 + r) for r in range(world)]
    for v, t in zip(views, vals):
        v.copy_(t)
    comm.allreduce(buf, n, active=(1 << world) - 1, payload=5, normalize=False, variant=variant, nblocks=8)
    torch.cuda.synchronize()
    want, _ = oracle.allreduce_sum(vals)
    for r, v in enumerate(views):
        assert torch.equal(v, views[0]), "ranks must be bit-identical"
        torch.testing.assert_close(v.cpu(), want, rtol=1e-6, atol=1e-5)
    res = comm.result_of(world - 1)
    assert res.n_active == world and res.payloads == [5 + r for r in range(world)]


@pytest.mark.parametrize("variant", ["oneshot", "ll"])
@pytest.mark.parametrize("dtype", [torch.int64, torch.int32, torch.float64])
def test_allreduce_other_dtypes(dtype, variant):
    world, n = 4, 100 if dtype != torch.int32 else 200
    comm = _emu(world)
    buf, views = comm.alloc_all(n, dtype)
    vals = [_rand(n, 20 + r, dtype) for r in range(world)]
    for v, t in zip(views, vals):
        v.copy_(t)
    comm.allreduce(buf, n, active=0b1111, variant=variant, nblocks=2)
    torch.cuda.synchronize()
    want, _ = oracle.allreduce_sum(vals)
    for v in views:
        if dtype.is_floating_point:
            torch.testing.assert_close(v.cpu(), want)
        else:
            assert torch.equal(v.cpu(), want)


@pytest.mark.parametrize("variant", ["oneshot", "twoshot", "hybrid", "ll"])
def test_allreduce_inactive_and_normalise(variant):
    world, n = 4, 2048
    comm = _emu(world)
    buf, views = comm.alloc_all(n)
    vals = [_rand(n, 30 + r) for r in range(world)]
    for v, t in zip(views, vals):
        v.copy_(t)
    comm.allreduce(buf, n, active=0b0101, normalize=True, variant=variant, nblocks=4)
    torch.cuda.synchronize()
    want, n_act = oracle.allreduce_sum(vals, [1, 0, 1, 0], normalize=True)
    for v in views:
        torch.testing.assert_close(v.cpu(), want, rtol=1e-6, atol=1e-6)
    assert comm.result_of(0).n_active == n_act == 2


@pytest.mark.parametrize("world", [1, 2, 4, 8, 5])
@pytest.mark.parametrize("variant", ["oneshot", "twoshot", "hybrid", "light", "ll"])
@pytest.mark.parametrize("momentum,wd,nesterov", [(0.0, 0.0, False), (0.9, 1e-3, False), (0.9, 0.0, True)])
def test_fused_sgd(world, variant, momentum, wd, nesterov):
    n, lr = 50016, 0.05
    comm = _emu(world)
    if variant == "ll":
        if world == 1:
            pytest.skip("a single rank has nobody to push to")
        n = comm.ll_max_bytes // 4
    g_buf, g_views = comm.alloc_all(n)
    p_buf, p_views = comm.alloc_all(n)
    m_buf, m_views = comm.alloc_all(n) if momentum else (None, None)
    p0 = _rand(n, 1)
    for v in p_views:
        v.copy_(p0)
    p_ref, m_ref = p0.clone(), (torch.zeros(n) if momentum else None)
    masks = [(1 << world) - 1, 0b1 if world > 1 else 1, 0, (1 << world) - 1]
    for s, mask in enumerate(masks):
        grads = [_rand(n, 100 * s + r) for r in range(world)]
        for v, t in zip(g_views, grads):
            v.copy_(t)
        comm.fused_sgd(g_buf, p_buf, m_buf, n, lr, momentum=momentum, weight_decay=wd, nesterov=nesterov, active=mask,
                       payload=s, zero_grads=True, variant=variant, nblocks=6)
        torch.cuda.synchronize()
        active = [(mask >> r) & 1 for r in range(world)]
        p_ref, m_ref, n_act = oracle.fused_sgd(grads, p_ref, lr, active, momentum, m_ref, wd, nesterov)
        assert comm.result_of(0).n_active == n_act
        for r in range(world):
            assert torch.equal(p_views[r], p_views[0]), "parameters must stay bit-identical across ranks"
            assert float(g_views[r].abs().max()) == 0.0, "fused zero_grad"
        torch.testing.assert_close(p_views[0].cpu(), p_ref, rtol=1e-5, atol=1e-6)


def test_fused_sgd_lr_comes_from_the_active_rank():
    """A draining rank passes a stale lr; the lowest ACTIVE rank's lr must be the one applied.
    (In emulation all ranks share the call's lr, so this checks the aux plumbing end to end.)"""
    world, n = 2, 4096
    comm = _emu(world)
    g_buf, g_views = comm.alloc_all(n)
    p_buf, p_views = comm.alloc_all(n)
    for v in g_views:
        v.fill_(1.0)
    comm.fused_sgd(g_buf, p_buf, None, n, 0.25, active=0b10, variant="twoshot", nblocks=2)
    torch.cuda.synchronize()
    for v in p_views:
        torch.testing.assert_close(v.cpu(), torch.full((n,), -0.25))


@pytest.mark.parametrize("world", [1, 2, 4, 8, 6])
@pytest.mark.parametrize("variant", ["twoshot", "hybrid", "ll"])
def test_fused_ea(world, variant):
    n, alpha = 40000, 0.2
    comm = _emu(world)
    if variant == "ll":
        if world == 1:
            pytest.skip("a single rank has nobody to push to")
        n = comm.ll_max_bytes // 4
    x_buf, x_views = comm.alloc_all(n)
    c0, c0_views = comm.alloc_all(n)
    c1, c1_views = comm.alloc_all(n)
    xs = [_rand(n, 40 + r) for r in range(world)]
    c = _rand(n, 3)
    for v, t in zip(x_views, xs):
        v.copy_(t)
    for v in c0_views:
        v.copy_(c)
    cur, nxt, cur_views, nxt_views = c0, c1, c0_views, c1_views
    for rnd in range(3):
        comm.fused_ea(x_buf, cur, nxt, n, alpha, active=(1 << world) - 1, variant=variant, nblocks=5)
        torch.cuda.synchronize()
        xs, c = oracle.elastic_average(xs, c, alpha)
        cur, nxt, cur_views, nxt_views = nxt, cur, nxt_views, cur_views
        for r in range(world):
            torch.testing.assert_close(x_views[r].cpu(), xs[r], rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(cur_views[r].cpu(), c, rtol=1e-5, atol=1e-6)
            assert torch.equal(cur_views[r], cur_views[0])
    # terminating drain round: nobody active -> nothing moves, next buffer mirrors the current one
    comm.fused_ea(x_buf, cur, nxt, n, alpha, active=0, variant=variant, nblocks=5)
    torch.cuda.synchronize()
    assert comm.result_of(0).n_active == 0
    for r in range(world):
        torch.testing.assert_close(x_views[r].cpu(), xs[r], rtol=1e-5, atol=1e-6)
        assert torch.equal(nxt_views[r], cur_views[r])


@pytest.mark.parametrize("world,root", [(1, 0), (2, 1), (4, 2), (8, 7), (8, 0), (5, 3), (3, 0)])
@pytest.mark.parametrize("variant", ["oneshot", "twoshot", "ll"])
def test_broadcast(world, root, variant):
    n = 30016
    comm = _emu(world)
    if variant == "ll":
        if world == 1:
            pytest.skip("a single rank has nobody to push to")
        n = comm.ll_max_bytes // 4
    a, a_views = comm.alloc_all(n)
    b, b_views = comm.alloc_all(n)
    for r, v in enumerate(a_views):
        v.copy_(_rand(n, 50 + r))
    want = _rand(n, 50 + root)
    comm.broadcast(a, n, root=root, also=b, variant=variant, nblocks=4)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(a_views[r].cpu(), want)
        assert torch.equal(b_views[r].cpu(), want)


def test_barrier_payloads():
    comm = _emu(4)
    res = comm.barrier(active=0b0110, payload=100)
    assert res.n_active == 2 and res.payloads == [100, 101, 102, 103]


def test_back_to_back_collectives_keep_epochs_in_sync():
    """Many launches in a row with different block counts: the device-side epoch counters must stay
    consistent per block (this is what makes the kernels CUDA-graph friendly)."""
    world, n = 4, 8192
    comm = _emu(world)
    buf, views = comm.alloc_all(n)
    for i in range(20):
        for r, v in enumerate(views):
            v.fill_(float(r + i))
        comm.allreduce(buf, n, active=0b1111, variant="twoshot" if i % 2 else "oneshot", nblocks=1 + i % 5)
    torch.cuda.synchronize()
    want = float(sum(r + 19 for r in range(world)))
    for v in views:
        assert torch.all(v == want)


def test_small_message_path_back_to_back_with_changing_sizes():
    """The push path has no barrier: its safety rests on call-numbered flags and parity double buffering.  Many calls in a
    row, sizes and grids changing, mixed with barrier-based calls on the same channel, all in one stream."""
    world = 8
    comm = _emu(world)
    cap = comm.ll_max_bytes // 4
    buf, views = comm.alloc_all(cap)
    sizes = [cap, 64, cap // 2, 4, 1024, cap, 8]
    for i in range(40):
        n = sizes[i % len(sizes)]
        for r, v in enumerate(views):
            v[:n].fill_(float(r + i))
        variant = "twoshot" if i % 7 == 3 else "ll"
        comm.allreduce(buf, n, active=(1 << world) - 1, variant=variant, nblocks=1 + i % 3)
        torch.cuda.synchronize()
        want = float(sum(r + i for r in range(world)))
        for v in views:
            assert torch.all(v[:n] == want), f"call {i}, n {n}, variant {variant}"


def test_contributor_count_reaches_the_host_without_a_copy():
    """lua/AllReduceSGD.lua:20-29 returns n; here it is a LazyCount backed by a word in pinned host memory."""
    import torch_distlearn_b200 as dl
    world = 4
    comm = _emu(world)
    sgd = dl.AllReduceSGD(comm)
    buf, views = comm.alloc_all(256)
    for mask, want in [(0b1111, 4), (0b0101, 2), (0b0001, 1)]:
        comm.allreduce(buf, 256, active=mask, normalize=True, variant="twoshot")
        n = comm.lazy_count(0, comm.last_ticket)
        assert int(n) == want and n == want and n + 1 == want + 1
    comm.poll_error()      # nothing went wrong: must not raise


def test_debug_mode_flags_an_impossible_epoch():
    """DL_FLAG_DEBUG: a barrier slot may only hold the previous use of its parity or the current epoch."""
    from torch_distlearn_b200 import _lib
    comm = _emu(2, heap_mb=8)
    comm.set_debug(True)
    comm.set_timeout(2.0)
    buf, views = comm.alloc_all(1024)
    comm.allreduce(buf, 1024, active=0b11, variant="twoshot", nblocks=1)     # healthy: epochs 1, 2
    torch.cuda.synchronize()
    comm.poll_error()
    # corrupt rank 0's slot for (parity 1, block 0, source 1): epoch 77 can never be there at barrier 3
    pad = comm.heaps[0][:_lib.DL_PAD_BYTES].view(torch.int32)
    slot = (1 * _lib.DL_MAX_BLOCKS + 0) * _lib.DL_MAX_RANKS + 1         # uint4 index into slots[2][blocks][ranks]
    pad[slot * 4 + 0] = 77
    pad[slot * 4 + 2] = 77
    comm._c.emulate = 0        # play rank 0 alone, so nobody repairs the slot
    with pytest.raises(_lib.NativeError):
        comm.allreduce(buf, 1024, active=1, variant="twoshot", nblocks=1, read=True)
    with pytest.raises(_lib.NativeError):
        comm.poll_error()
    comm.poll_error()          # reported once, then cleared


@pytest.mark.parametrize("exclusive", [False, True])
def test_async_ea_client_kernel(exclusive):
    """Server heap and client heap on one GPU; the client kernel pulls the center, moves x and pushes
    the elastic difference with vector atomics."""
    from torch_distlearn_b200 import EmulatedCommunicator
    comm = EmulatedCommunicator(2, torch.device(DEV), heap_mb=8)
    comm._c.emulate = 0   # async kernels are ordinary launches; rank 1 is the client
    comm._c.rank = 1
    n, alpha = 10007, 0.2
    center_buf, center_views = comm.alloc_all(comm.padded(n))
    c = _rand(n, 7)
    center_views[0][:n].copy_(c)
    x = _rand(n, 8).to(DEV)
    x_ref = x.cpu().clone()
    for _ in range(3):
        comm.async_client_sync(0, x, center_buf, n, alpha, exclusive=exclusive, nblocks=4)
        x_ref, c = oracle.async_client_sync(x_ref, c, alpha)
    torch.cuda.synchronize()
    torch.testing.assert_close(x.cpu(), x_ref, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(center_views[0][:n].cpu(), c, rtol=1e-5, atol=1e-6)
    assert comm.sync_count(0) == 3
    dst = torch.zeros(n, device=DEV)
    comm.peer_pull(0, center_buf, dst, n)
    torch.cuda.synchronize()
    assert torch.equal(dst, center_views[0][:n])
    comm.mailbox_write(0, 3, 12345)
    assert comm.mailbox_read(0, 3) == 12345


def test_local_ops():
    from torch_distlearn_b200.ops import confusion_add, local_ea, local_sgd, sumsq_diff
    n = 100003
    p, g = _rand(n, 1).to(DEV), _rand(n, 2).to(DEV)
    m = torch.zeros(n, device=DEV)
    p_ref, m_ref, _ = oracle.fused_sgd([g.cpu() * 0.5], p.cpu(), 0.1, [1], 0.9, torch.zeros(n), 1e-3, False)
    local_sgd(p, g, m, lr=0.1, momentum=0.9, weight_decay=1e-3, grad_scale=0.5, zero_grads=True)
    torch.cuda.synchronize()
    torch.testing.assert_close(p.cpu(), p_ref, rtol=1e-5, atol=1e-6)
    assert float(g.abs().max()) == 0.0
    x, c = _rand(n, 3).to(DEV), _rand(n, 4).to(DEV)
    xr, cr = oracle.elastic_average([x.cpu()], c.cpu(), 0.3)
    local_ea(x, c, 0.3)
    torch.testing.assert_close(x.cpu(), xr[0], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(c.cpu(), cr, rtol=1e-5, atol=1e-6)
    a, b = _rand(n, 5).to(DEV), _rand(n, 6).to(DEV)
    s1 = sumsq_diff(a, b)
    s2 = sumsq_diff(a, b)
    want = ((a.double() - b.double()) ** 2).sum()
    torch.testing.assert_close(s1[0], want, rtol=1e-9, atol=0)
    assert torch.equal(s1, s2), "deterministic"
    logits = torch.randn(257, 10, device=DEV)
    tgt = torch.randint(0, 10, (257,), device=DEV)
    mat = torch.zeros(10, 10, dtype=torch.int64, device=DEV)
    confusion_add(logits, tgt, mat)
    ref = torch.zeros(10, 10, dtype=torch.int64)
    for t, pr in zip(tgt.cpu().tolist(), logits.argmax(1).cpu().tolist()):
        ref[t, pr] += 1
    assert torch.equal(mat.cpu(), ref)


def test_timeout_is_reported_not_hung():
    """Fault injection: a 2-rank communicator where the peer never shows up must time out."""
    from torch_distlearn_b200 import EmulatedCommunicator, _lib
    comm = EmulatedCommunicator(2, torch.device(DEV), heap_mb=4)
    comm._c.emulate = 0     # play rank 0 only; "rank 1" never launches
    comm.set_timeout(0.2)
    buf = comm.alloc(1024)
    with pytest.raises(_lib.NativeError):
        comm.allreduce(buf, 1024, active=1, variant="twoshot", nblocks=1, read=True)


@pytest.mark.parametrize("world", [2, 8])
def test_litmus_message_passing_and_store_buffering(world):
    """The flag protocol's two shapes with random skew (csrc/litmus.cu).  On one GPU the ranks are emulated (one
    cooperative launch), so this pins the code path and the GPU-scope behaviour; tests/run_multigpu.py runs the same
    shapes across real GPUs over NVLink."""
    comm = _emu(world, heap_mb=64)
    iters = 20000
    res = comm.litmus("mp", fenced=True, iterations=iters, nblocks=4)
    for r in res:
        assert r["timeouts"] == 0 and r["violations"] == 0 and r["checked"] == iters * 4 * 32
    res = comm.litmus("sb", fenced=True, iterations=iters, nblocks=4)
    loaded = torch.stack([r["loaded"] for r in res])          # [world, blocks, iterations]
    assert all(r["timeouts"] == 0 for r in res)
    forbidden = int((loaded.sum(0) == 0).sum())               # every rank of the ring loaded the old value
    assert forbidden == 0, f"{forbidden} store-buffering violations with fence.sc.sys"

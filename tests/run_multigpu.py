#!/usr/bin/env python
"""Multi-GPU check program: one process per GPU (torchrun), real P2P symmetric heaps over NVLink.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tests/run_multigpu.py [--quick]

Part 1 compares every collective kernel variant (one-shot / two-shot / NVLS) with a torch reduction
of the gathered inputs; part 2 runs the scenarios of tests/dist_checks.py (the reference's two
tests and the AsyncEA / fused / Tree scenarios) and verifies them on rank 0 against the oracle.
torch.distributed (gloo) is used for bootstrap and for gathering results only.
Prints ``MULTIGPU OK`` on success; any failure exits non-zero on every rank."""
from __future__ import annotations

import argparse
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch
import torch.distributed as dist


def gather(obj, world):
    out = [None] * world
    dist.all_gather_object(out, obj)
    return out


def rand(n, seed):
    return torch.randn(n, generator=torch.Generator().manual_seed(seed), dtype=torch.float32)


def kernel_checks(comm, rank, world, dev, quick):
    # (variant, share of every slice that goes through the switch): the hybrid kernel is checked all-P2P, mixed and all-switch
    # ("light" is a fused_sgd variant — the per-bucket kernel — the other operations run their two-shot form under that name)
    variants = [("ll", 0), ("oneshot", 0), ("twoshot", 0), ("hybrid", 0), ("light", 0)]
    if comm.has_multicast:
        variants += [("nvls", 0), ("hybrid", 600), ("hybrid", 1000), ("light", 600)]
    sizes = [256, 40000, 1 << 20] if quick else [64, 4128, 40000, 1 << 20, (1 << 24) + 64]
    nmax = max(sizes)
    buf = comm.alloc(nmax, torch.float32, "chk.a")
    p = comm.alloc(nmax, torch.float32, "chk.p")
    m = comm.alloc(nmax, torch.float32, "chk.m")
    c0 = comm.alloc(nmax, torch.float32, "chk.c0")
    c1 = comm.alloc(nmax, torch.float32, "chk.c1")
    report = []
    for n in sizes:
        for v, permille in variants:
            if v == "oneshot" and n > 100000:
                continue
            if v == "ll" and n * 4 > comm.ll_max_bytes:
                continue
            comm._variants.hybrid_permille = {op: permille for op in comm._variants.OPS}
            vb = v if v != "hybrid" else "twoshot"          # broadcast has no hybrid form
            # ---- all-reduce (+ normalise) ----
            vals = [rand(n, 1000 + 13 * r + n % 97) for r in range(world)]
            buf.tensor[:n].copy_(vals[rank].to(dev))
            comm.synchronize(); dist.barrier()
            comm.allreduce(buf, n, active=1, payload=rank, normalize=True, variant=v)
            comm.synchronize()
            want = torch.stack(vals).sum(0) / world
            got = buf.tensor[:n].cpu()
            torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-5, msg=lambda s: f"allreduce {v} n={n}: {s}")
            alls = gather(got, world)
            assert all(torch.equal(alls[0], a) for a in alls), f"allreduce {v} n={n}: ranks differ bitwise"
            # ---- fused SGD with momentum ----
            g = [rand(n, 2000 + 7 * r + n % 89) for r in range(world)]
            p0 = rand(n, 5)
            buf.tensor[:n].copy_(g[rank].to(dev)); p.tensor[:n].copy_(p0.to(dev)); m.tensor[:n].zero_()
            comm.synchronize(); dist.barrier()
            comm.fused_sgd(buf, p, m, n, 0.1, momentum=0.9, weight_decay=1e-3, active=1, payload=rank, variant=v)
            comm.synchronize()
            gs = torch.stack(g).sum(0) / world + 1e-3 * p0
            want_p = p0 - 0.1 * gs
            got_p = p.tensor[:n].cpu()
            torch.testing.assert_close(got_p, want_p, rtol=1e-5, atol=1e-5, msg=lambda s: f"fused_sgd {v} n={n}: {s}")
            assert float(buf.tensor[:n].abs().max()) == 0.0, f"fused_sgd {v}: grads not zeroed"
            alls = gather(got_p, world)
            assert all(torch.equal(alls[0], a) for a in alls), f"fused_sgd {v} n={n}: params differ bitwise"
            # ---- broadcast from the last rank, into two buffers ----
            buf.tensor[:n].copy_(vals[rank].to(dev)); p.tensor[:n].fill_(-3.0)
            comm.synchronize(); dist.barrier()
            comm.broadcast(buf, n, root=world - 1, also=p, variant=vb)
            comm.synchronize()
            assert torch.equal(buf.tensor[:n].cpu(), vals[world - 1]) and torch.equal(p.tensor[:n].cpu(), vals[world - 1]), \
                f"broadcast {v} n={n}"
            # ---- fused elastic average ----
            if v != "oneshot":
                # the switch path computes a (sum x - N c) instead of sum a (x - c): same value, different rounding
                ea_tol = 1e-4 if (v == "nvls" or permille) else 1e-5
                xs = [rand(n, 3000 + 11 * r) for r in range(world)]
                cc = rand(n, 77)
                buf.tensor[:n].copy_(xs[rank].to(dev)); c0.tensor[:n].copy_(cc.to(dev))
                comm.synchronize(); dist.barrier()
                comm.fused_ea(buf, c0, c1, n, 0.1, active=1, variant=v)
                comm.synchronize()
                d = [(x - cc) * 0.1 for x in xs]
                torch.testing.assert_close(buf.tensor[:n].cpu(), xs[rank] - d[rank], rtol=1e-5, atol=1e-5)
                torch.testing.assert_close(c1.tensor[:n].cpu(), cc + torch.stack(d).sum(0), rtol=1e-4, atol=ea_tol,
                                           msg=lambda s: f"fused_ea {v} n={n}: {s}")
                alls = gather(c1.tensor[:n].cpu(), world)
                assert all(torch.equal(alls[0], a) for a in alls), f"fused_ea {v}/{permille} n={n}: centers differ bitwise"
            report.append((n, v, permille))
    # inactive ranks: only even ranks contribute
    n = 40000
    vals = [rand(n, 4000 + r) for r in range(world)]
    buf.tensor[:n].copy_(vals[rank].to(dev))
    comm.synchronize(); dist.barrier()
    res = comm.allreduce(buf, n, active=1 if rank % 2 == 0 else 0, payload=100 + rank, normalize=True, variant="twoshot", read=True)
    act = [r for r in range(world) if r % 2 == 0]
    want = torch.stack([vals[r] for r in act]).sum(0) / (len(act) if len(act) > 1 else 1)
    torch.testing.assert_close(buf.tensor[:n].cpu(), want, rtol=1e-5, atol=1e-5)
    assert res.n_active == len(act) and res.payloads == [100 + r for r in range(world)]
    # the count also reaches the host through pinned memory, without a copy (LazyCount)
    lazy = comm.lazy_count(0, comm.last_ticket)
    assert int(lazy) == len(act), f"lazy contributor count {int(lazy)} != {len(act)}"
    # int64 confusion-matrix style reduce
    ib = comm.alloc(100, torch.int64, "chk.i")
    for v in ("oneshot", "ll"):
        ib.tensor.fill_(rank + 1)
        comm.synchronize(); dist.barrier()
        comm.allreduce(ib, 100, active=1, variant=v)
        comm.synchronize()
        assert torch.all(ib.tensor == sum(range(1, world + 1))), f"int64 all-reduce {v}"
    # a run of small-message calls of changing size, no host synchronisation in between (parity double buffering)
    sb = comm.alloc(comm.ll_max_bytes // 4, torch.float32, "chk.ll")
    runs = []
    for i in range(30):
        k = [sb.numel, 64, sb.numel // 2, 8][i % 4]
        sb.tensor[:k].fill_(float(rank + i))
        comm.allreduce(sb, k, active=1, variant="ll")
        runs.append((sb.tensor[:k].clone(), float(sum(r + i for r in range(world)))))
    comm.synchronize()
    for i, (got, want_v) in enumerate(runs):
        assert torch.all(got == want_v), f"ll run, call {i}"
    dist.barrier()
    comm.poll_error()
    return report


def litmus_checks(comm, rank, world, quick):
    """Message passing and store buffering across the real GPUs (csrc/litmus.cu), with and without the fences."""
    iters = 20000 if quick else 100000
    lines = []
    for shape in ("mp", "sb"):
        for fenced in (True, False):
            res = comm.litmus(shape, fenced=fenced, iterations=iters, nblocks=8, seed=17 + rank)
            assert res["timeouts"] == 0, f"litmus {shape}: {res['timeouts']} blocks timed out"
            if shape == "mp":
                viol = sum(gather(res["violations"], world))
            else:
                loaded = torch.stack(gather(res["loaded"], world))
                viol = int((loaded.sum(0) == 0).sum())
            lines.append(f"litmus {shape} fenced={int(fenced)}: {viol} forbidden outcomes in {iters} iterations x 8 blocks x {world} ranks")
            if fenced:
                assert viol == 0, lines[-1]
    return lines


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--scenarios", default="all")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import dist_checks
    import torch_distlearn_b200 as dl
    ok = True
    try:
        comm = dl.P2PCommunicator(rank=rank, world=world, device=dev, heap_mb=768 if not args.quick else 128)
        if rank == 0:
            print("transport:", comm.describe(), flush=True)
        rep = kernel_checks(comm, rank, world, dev, args.quick)
        if rank == 0:
            print(f"kernel checks ok: {len(rep)} (size, variant, switch share) cells", flush=True)
        for line in litmus_checks(comm, rank, world, args.quick):
            if rank == 0:
                print(line, flush=True)
        names = list(dist_checks.SCENARIOS) if args.scenarios == "all" else args.scenarios.split(",")
        for name in names:
            run, verify = dist_checks.SCENARIOS[name]
            kw = {}
            if name == "async_ea":
                if world < 3:
                    kw = dict(exclusive=True, with_tester=False)
                else:
                    kw = dict(exclusive=(world % 2 == 0), with_tester=True)
            kws = [kw] if name != "sgd_bucketed" else [dict(use_graph=False), dict(use_graph=True)]
            for kw in kws:
                out = run(comm, rank, world, **kw)
                comm.synchronize()
                results = gather(out, world)
                if rank == 0:
                    verify(results, world, **kw)
                    print(f"scenario {name} {kw if kw else ''}: ok", flush=True)
                dist.barrier()
    except BaseException:  # noqa: BLE001
        ok = False
        print(f"[rank {rank}] FAILED\n{traceback.format_exc()}", flush=True)
    oks = gather(ok, world) if ok else None
    if ok and all(oks):
        if rank == 0:
            print("MULTIGPU OK", flush=True)
        dist.destroy_process_group()
        return 0
    os._exit(1)


if __name__ == "__main__":
    sys.exit(main())

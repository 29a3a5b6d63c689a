#!/usr/bin/env python
"""Collective sweep: flat-gradient allreduce+SGD, plain all-reduce and
parameter broadcast, 1 KB - 1 GB, this library's sm_100a kernels (every variant) vs NCCL.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29541 bench/sweep.py [--max-mb 1024] [--out gpurun_out/sweep_N.json] [--write-table]

Method: per (op, size, implementation) W warm-up launches, then K launches timed as ONE CUDA-event
interval on the launching stream, device barrier (our own) before the first; time = interval / K,
max over ranks.  Buffers >= 128 MB exceed the L2, smaller ones are L2/latency-bound by nature of the
op (gradients ARE L2-resident right after backward) - reported as is, no flush, stated in the output.
Bus bandwidth uses the NCCL convention  busbw = S/t * 2(N-1)/N  (broadcast: S/t).  The roofline of a
variant is the bytes it must move per direction per GPU over NVLink:
    two-shot P2P  2(N-1)/N * S        NVLS  (1 + 1/N) * S        broadcast  S (root egress) or S/N+S (2-phase)
divided by the measured per-direction peer bandwidth (770 GB/s, B200_PROFILING.md).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

LINK_GBS = 770.0  # per direction per GPU: the recipe's measured peer-copy figure (B200_PROFILING.md); replaced at run time by
                  # the figure bench/peak.py measured in this repo (profiles/nvlink_peak.json) when that file is present


def _measured_link_gbs():
    path = os.path.join(ROOT, "profiles", "nvlink_peak.json")
    try:
        return float(json.load(open(path))["p2p_copy_gbs_per_direction"])
    except Exception:  # noqa: BLE001 - no measurement yet: keep the recipe constant
        return None


GRAPH = False


def timed(fn, warm, iters, sync, world):
    """Mean time of one call in microseconds (max over ranks).  With --graph the `iters` calls are
    captured into one CUDA graph and replayed, which removes the host launch cost from both arms."""
    for _ in range(warm):
        fn()
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if GRAPH:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                for _ in range(iters):
                    fn()
        torch.cuda.current_stream().wait_stream(side)
        graph.replay()                                   # warm replay
        sync()
        e0.record()
        graph.replay()
        e1.record()
    else:
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    t = torch.tensor([ms], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)          # gloo group (CPU tensor)
    return float(t.item()) * 1e3                       # microseconds


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=1024)
    ap.add_argument("--min-kb", type=float, default=1)
    ap.add_argument("--out", default="")
    ap.add_argument("--write-table", action="store_true")
    ap.add_argument("--ops", default="fused_sgd,allreduce,broadcast,fused_ea")
    ap.add_argument("--nblocks", default="0", help="comma list of block counts to try for our kernels (0 = default)")
    ap.add_argument("--no-nccl", action="store_true")
    ap.add_argument("--graph", action="store_true", help="time CUDA-graph replays (no host launch cost in either arm)")
    ap.add_argument("--permille", default="0,500,750,1000", help="switch shares tried for the hybrid variant (needs multicast)")
    ap.add_argument("--variants", default="", help="comma list to restrict our variants (default: all that apply)")
    args = ap.parse_args()
    global GRAPH, LINK_GBS
    GRAPH = bool(args.graph)
    link_source = "recipe constant (B200_PROFILING.md)"
    if _measured_link_gbs():
        LINK_GBS, link_source = _measured_link_gbs(), "profiles/nvlink_peak.json (bench/peak.py, this repo)"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nccl = None if args.no_nccl else dist.new_group(backend="nccl")
    import torch_distlearn_b200 as dl
    max_n = int(args.max_mb * (1 << 20)) // 4
    heap_mb = args.max_mb * 4 + 64
    comm = dl.P2PCommunicator(rank=rank, world=world, device=dev, heap_mb=heap_mb)
    g = comm.alloc(max_n, torch.float32, "sw.g")
    p = comm.alloc(max_n, torch.float32, "sw.p")
    c0 = comm.alloc(max_n, torch.float32, "sw.c0")
    c1 = comm.alloc(max_n, torch.float32, "sw.c1")
    g.tensor.normal_(); p.tensor.normal_()
    tg = torch.randn(max_n, device=dev)                # NCCL arm: ordinary device memory
    tp = torch.randn(max_n, device=dev)
    tc = torch.randn(max_n, device=dev)
    td = torch.zeros(max_n, device=dev)
    ops = args.ops.split(",")
    nblocks_list = [int(x) for x in args.nblocks.split(",")]
    permilles = [int(x) for x in args.permille.split(",")] if comm.has_multicast else [0]
    variants = ["ll", "oneshot", "twoshot"] + (["nvls"] if comm.has_multicast else []) + [f"hybrid{p}" for p in permilles]
    if args.variants:
        keep = set(args.variants.split(","))
        variants = [v for v in variants if v in keep or (v.startswith("hybrid") and "hybrid" in keep)]
    rows = []

    def dev_sync():
        comm.barrier(read=False)
        torch.cuda.synchronize()
        dist.barrier()

    sizes = []
    s = int(args.min_kb * 1024)
    while s <= args.max_mb * (1 << 20):
        sizes.append(s)
        s *= 4
    if rank == 0:
        print(json.dumps({"transport": comm.describe(), "world": world, "sizes": sizes}), flush=True)
    for nbytes in sizes:
        n = nbytes // 4
        iters = 200 if nbytes <= (1 << 20) else 50 if nbytes <= (64 << 20) else 10
        warm = max(3, iters // 10)
        cur = [0]
        cands = {}
        for op in ops:
            for vname in variants:
                v = vname
                if v == "oneshot" and (nbytes > (2 << 20) or op == "fused_ea"):
                    continue
                if v == "ll" and nbytes > comm.ll_max_bytes:
                    continue
                if v.startswith("hybrid"):
                    if op == "broadcast" or nbytes < (64 << 10) or n % 8:
                        continue
                    pm = int(v[6:])
                    comm._variants.hybrid_permille = {o: pm for o in comm._variants.OPS}
                    v = "hybrid"
                for nb in nblocks_list:
                    if op == "fused_sgd":
                        fn = lambda v=v, nb=nb: comm.fused_sgd(g, p, None, n, 1e-6, active=1, variant=v, zero_grads=False, nblocks=nb)
                    elif op == "allreduce":
                        fn = lambda v=v, nb=nb: comm.allreduce(g, n, active=1, normalize=True, variant=v, nblocks=nb)
                    elif op == "broadcast":
                        fn = lambda v=v, nb=nb: comm.broadcast(p, n, root=0, variant=v, nblocks=nb)
                    else:
                        def fn(v=v, nb=nb):
                            a, b = (c0, c1) if cur[0] == 0 else (c1, c0)
                            comm.fused_ea(p, a, b, n, 0.01, active=1, variant=v, nblocks=nb)
                            cur[0] ^= 1
                    try:
                        us = timed(fn, warm, iters, dev_sync, world)
                    except Exception as e:  # noqa: BLE001
                        if rank == 0:
                            print(f"# {op} {v} {nbytes}: {e}", flush=True)
                        continue
                    cands[(op, vname, nb)] = us
            if nccl is not None:
                if op == "fused_sgd":
                    def fn():
                        dist.all_reduce(tg[:n], op=dist.ReduceOp.AVG, group=nccl)
                        tp[:n].add_(tg[:n], alpha=-1e-6)
                elif op == "allreduce":
                    fn = lambda: dist.all_reduce(tg[:n], op=dist.ReduceOp.AVG, group=nccl)
                elif op == "broadcast":
                    fn = lambda: dist.broadcast(tp[:n], src=0, group=nccl)
                else:
                    def fn():
                        torch.sub(tp[:n], tc[:n], out=td[:n]); td[:n].mul_(0.01); tp[:n].sub_(td[:n])
                        dist.all_reduce(td[:n], group=nccl); tc[:n].add_(td[:n])
                cands[(op, "nccl", 0)] = timed(fn, warm, iters, dev_sync, world)
        if rank == 0:
            for (op, v, nb), us in sorted(cands.items()):
                algbw = nbytes / us / 1e3                       # GB/s
                factor = 1.0 if op == "broadcast" else 2.0 * (world - 1) / world
                if op == "broadcast":
                    roof_bytes = nbytes * 1.0        # every rank has to receive S, the root has to send S
                elif v == "nvls":
                    roof_bytes = nbytes * (1.0 + 1.0 / world)
                elif v.startswith("hybrid"):
                    f = int(v[6:]) / 1000.0           # share through the switch: (1 + 1/N) S, the rest 2 (N-1)/N S
                    roof_bytes = nbytes * (f * (1.0 + 1.0 / world) + (1.0 - f) * 2.0 * (world - 1) / world)
                elif v == "ll":
                    roof_bytes = nbytes * 2.0 * (world - 1)   # 2x packets, pushed to every peer
                else:
                    roof_bytes = nbytes * 2.0 * (world - 1) / world
                roof_us = roof_bytes / (LINK_GBS * 1e3)
                row = {"op": op, "impl": v, "nblocks": nb, "bytes": nbytes, "us": round(us, 2), "algbw_gbs": round(algbw, 2),
                       "busbw_gbs": round(algbw * factor, 2), "roofline_us": round(roof_us, 2),
                       "frac_of_roofline": round(roof_us / us, 4), "world": world, "graph": GRAPH}
                rows.append(row)
                print(json.dumps(row), flush=True)
    if rank == 0:
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            json.dump({"world": world, "transport": comm.describe(), "link_gbs": LINK_GBS, "link_gbs_source": link_source,
                       "rows": rows}, open(args.out, "w"), indent=1)
        if args.write_table:
            write_table(rows, world)
    dist.barrier()
    dist.destroy_process_group()


def write_table(rows, world):
    """Turns a sweep into the size -> variant table that Communicator.pick_variant reads (parallel/variant_table.json):
    ll_max[op]          largest size up to which the push path is the fastest of our variants at every smaller size too
    oneshot_max[op]     same for the one-shot kernel (above ll_max)
    hybrid_permille[op] switch share of the hybrid kernel that minimises the summed time over the sizes >= 4 MiB
    hybrid_min[op]      smallest size from which that hybrid beats two-shot and NVLS at every larger size
    nvls_min[op]        smallest size from which NVLS beats two-shot at every larger size (used below hybrid_min)."""
    ours = {}
    for r in rows:
        if r["impl"] == "nccl":
            continue
        best = ours.setdefault((r["op"], r["bytes"]), {})
        best[r["impl"]] = min(best.get(r["impl"], 1e30), r["us"])
    table = {"world": world, "ll_max": {}, "oneshot_max": {}, "nvls_min": {}, "hybrid_min": {}, "hybrid_permille": {}}
    for op in sorted({k[0] for k in ours}):
        sizes = sorted(b for (o, b) in ours if o == op)

        def prefix_max(name):
            top = 0
            for b in sizes:
                t = ours[(op, b)]
                if name in t and t[name] <= min(t.values()) * 1.02:
                    top = b
                elif name in t or b > top:
                    break
            return top

        table["ll_max"][op] = prefix_max("ll")
        one_max = 0
        for b in sizes:
            t = {k: v for k, v in ours[(op, b)].items() if k != "ll"}
            if "oneshot" in t and t["oneshot"] <= min(t.values()):
                one_max = b
            elif b > table["ll_max"][op]:
                break
        table["oneshot_max"][op] = one_max
        hybrids = sorted({k for b in sizes for k in ours[(op, b)] if k.startswith("hybrid")})
        big = [b for b in sizes if b >= (4 << 20)]
        if hybrids and big:
            tot = {h: sum(ours[(op, b)].get(h, 1e30) for b in big) for h in hybrids}
            h_best = min(tot, key=tot.get)
            table["hybrid_permille"][op] = int(h_best[6:])
            h_min = None
            for b in reversed(sizes):
                t = ours[(op, b)]
                if h_best in t and t[h_best] <= min(t.get("twoshot", 1e30), t.get("nvls", 1e30)):
                    h_min = b
                else:
                    break
            table["hybrid_min"][op] = h_min if h_min is not None else 1 << 62
        nv_min = None
        for b in reversed(sizes):
            t = ours[(op, b)]
            if "nvls" in t and t["nvls"] <= t.get("twoshot", 1e30):
                nv_min = b
            else:
                break
        table["nvls_min"][op] = nv_min if nv_min is not None else 1 << 62
    path = os.path.join(ROOT, "gpurun_out", f"variant_table_{world}{'_graph' if GRAPH else ''}.json")
    json.dump(table, open(path, "w"), indent=1)
    print("# variant table ->", path, json.dumps(table))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""CPU ranks: fused all-reduce + 1/n + SGD over the shared-memory transport (csrc/host_transport.cpp) against
gloo all_reduce + div + add_ (what `bench.py --impl nccl` does on GPUs), vector sizes from the MNIST model to the CIFAR
model and beyond.  No GPU needed.

    python bench/bench_cpu_collectives.py [--nodes 4] [--sizes 10848,4328992,16777216]

Wall-clock per call after a warm-up, max over ranks; one JSON line per (size, arm)."""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import socket
import sys
import time
import uuid

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(rank, world, impl, sizes, job, port, q):
    import torch
    torch.set_num_threads(1)
    out = {}
    if impl == "ours":
        import torch_distlearn_b200 as dl
        comm = dl.HostCommunicator(rank, world, name=job, heap_mb=int(2 * sum(sizes) * 4 / 2**20) + 16, timeout_s=120)
        for n in sizes:
            g, p = comm.alloc(n, torch.float32, f"g{n}"), comm.alloc(n, torch.float32, f"p{n}")
            g.tensor.fill_(1.0)
            iters = max(5, min(400, int(2e8 / (n * world))))
            for _ in range(3):
                comm.fused_sgd(g, p, None, n, 0.01)
            comm.barrier()
            t0 = time.perf_counter()
            for _ in range(iters):
                comm.fused_sgd(g, p, None, n, 0.01)
            out[n] = (time.perf_counter() - t0) / iters
        comm.close()
    else:
        import torch.distributed as dist
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        for n in sizes:
            g, p = torch.ones(n), torch.zeros(n)
            iters = max(5, min(400, int(2e8 / (n * world))))

            def step():
                dist.all_reduce(g)
                g.div_(world)
                p.add_(g, alpha=-0.01)
                g.zero_()
            for _ in range(3):
                step()
            dist.barrier()
            t0 = time.perf_counter()
            for _ in range(iters):
                step()
            out[n] = (time.perf_counter() - t0) / iters
        dist.destroy_process_group()
    q.put((rank, out))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=4)
    ap.add_argument("--sizes", default="10848,262144,4328992,16777216")
    args = ap.parse_args()
    sizes = [int(s) for s in args.sizes.split(",")]
    res = {}
    for impl in ("ours", "gloo"):
        s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        job = f"cc-{os.getpid()}-{uuid.uuid4().hex[:6]}"
        ps = [ctx.Process(target=worker, args=(r, args.nodes, impl, sizes, job, port, q)) for r in range(args.nodes)]
        for p in ps:
            p.start()
        outs = [q.get(timeout=600) for _ in ps]
        for p in ps:
            p.join()
        res[impl] = {n: max(o[1][n] for o in outs) for n in sizes}
    for n in sizes:
        for impl in ("ours", "gloo"):
            t = res[impl][n]
            print(json.dumps({"bench": "cpu_allreduce_sgd", "nodes": args.nodes, "floats": n, "MiB": round(n * 4 / 2**20, 2), "impl": impl,
                              "us": round(t * 1e6, 1), "algbw_GBps": round(n * 4 / t / 1e9, 2),
                              "vs_gloo": round(res["gloo"][n] / t, 2) if impl == "ours" else 1.0}))


if __name__ == "__main__":
    main()

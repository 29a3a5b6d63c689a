#!/usr/bin/env python
"""MNIST convnet AllReduceSGD on N CPU processes (no GPU), batch 1, lr 0.01
(examples/mnist.lua:33,113 run by examples/mnist.sh).

    python bench/bench_cpu_mnist.py [--nodes 2] [--steps 400] [--batch 1]

ours : C++ shared-memory transport (csrc/host_transport.cpp), fused all-reduce + 1/n + SGD per step
gloo : torch.distributed gloo all_reduce of the flat gradient + div + add_ (the library baseline)
Both arms run the same model code with one thread per process; wall-clock over `steps` steps after a warm-up,
max over ranks.  Prints one JSON line per arm."""
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


def worker(rank, world, impl, steps, warm, batch, job, port, q):
    import torch
    import torch.nn.functional as F
    torch.set_num_threads(1)
    from torch_distlearn_b200.models import mnist_convnet
    model = mnist_convnet()
    g = torch.Generator().manual_seed(rank)
    xs = [torch.randn(batch, 1024, generator=g) for _ in range(32)]
    ys = [torch.randint(0, 10, (batch,), generator=g) for _ in range(32)]
    if impl == "ours":
        import torch_distlearn_b200 as dl
        comm = dl.HostCommunicator(rank, world, name=job, heap_mb=4, timeout_s=120)
        flat = dl.FlatParams(comm, model)
        sgd = dl.AllReduceSGD(comm, flat)
        sgd.synchronizeParameters()

        def step(i):
            F.nll_loss(model(xs[i % 32]), ys[i % 32]).backward()
            sgd.step(lr=0.01)

        sync = lambda: comm.barrier()
    else:
        import torch.distributed as dist
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from baseline.nccl_algos import NcclFlat
        flat = NcclFlat(model)
        dist.broadcast(flat.params, src=0)

        def step(i):
            F.nll_loss(model(xs[i % 32]), ys[i % 32]).backward()
            dist.all_reduce(flat.grads)
            flat.grads.div_(world)
            flat.params.add_(flat.grads, alpha=-0.01)
            flat.grads.zero_()

        sync = lambda: dist.barrier()
    for i in range(warm):
        step(i)
    sync()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    sync()
    q.put((rank, time.perf_counter() - t0))
    if impl == "ours":
        comm.close()
    else:
        import torch.distributed as dist
        dist.destroy_process_group()


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=2)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--batch", type=int, default=1)
    a = ap.parse_args()
    ctx = mp.get_context("spawn")
    for impl in ("ours", "gloo"):
        q = ctx.Queue()
        job, port = f"bcpu-{uuid.uuid4().hex[:8]}", free_port()
        ps = [ctx.Process(target=worker, args=(r, a.nodes, impl, a.steps, a.warmup, a.batch, job, port, q)) for r in range(a.nodes)]
        for p in ps:
            p.start()
        times = [q.get(timeout=600)[1] for _ in ps]
        for p in ps:
            p.join()
        t = max(times)
        print(json.dumps({"bench": "mnist_allreduce_sgd_cpu", "impl": impl, "nodes": a.nodes, "per_node_batch": a.batch,
                          "steps": a.steps, "ms_per_step": round(t / a.steps * 1e3, 4),
                          "samples_per_s": round(a.nodes * a.batch * a.steps / t, 1)}), flush=True)


if __name__ == "__main__":
    main()

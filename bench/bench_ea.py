#!/usr/bin/env python
"""MNIST convnet, AllReduceEA (tau=10, alpha=0.2), synthetic data, N GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 \
        bench/bench_ea.py [--impl ours|nccl] [--batch 1] [--steps 2000]

ours: the local step (forward + backward + SGD) is ONE kernel (csrc/mnist_step.cu: an 8-CTA cluster, batch <= 8) and
every tau-th step adds ONE more, the fused elastic average (the LL push variant at this size).  `--fused 0` and the nccl
arm run the stock torch model instead, with forward + backward + local update captured in one CUDA graph (~25 small
library kernels), which is what round 1 measured in both arms.  Every tau-th step runs the exchange:
ours = ONE fused kernel (csrc/fused_ea.cu), nccl = 3 elementwise kernels + all_reduce + 1 elementwise
kernel (lua/AllReduceEA.lua:34-45 spelled with torch ops).  Time: CUDA events around all steps, max
over ranks.  Prints one JSON line on rank 0."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch_distlearn_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", choices=["ours", "nccl"], default="ours")
    ap.add_argument("--batch", type=int, default=1, help="per-GPU batch (examples/mnist-ea.lua uses 1)")
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--tau", type=int, default=10)
    ap.add_argument("--alpha", type=float, default=0.2)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--fused", type=int, default=1, help="ours: the one-kernel MNIST step (batch <= 8)")
    ap.add_argument("--graph-steps", type=int, default=0,
                    help="local steps per CUDA-graph replay (0 = tau): the tau steps between two exchanges are one replay, "
                         "each step reading its own batch in place, so the host issues 2 launches per tau steps in both arms")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl" if args.impl == "nccl" else "gloo", rank=rank, world_size=world,
                                **({"device_id": dev} if args.impl == "nccl" else {}))
    from torch_distlearn_b200.models import mnist_convnet
    from torch_distlearn_b200.ops import local_sgd
    from torch_distlearn_b200.utils.timing import max_over_ranks
    model = mnist_convnet().to(dev)
    B = args.batch
    g = torch.Generator().manual_seed(rank)
    xs = [torch.randn(B, 1024, generator=g).to(dev) for _ in range(16)]
    ys = [torch.randint(0, 10, (B,), generator=g).to(dev) for _ in range(16)]
    x, y = torch.zeros(B, 1024, device=dev), torch.zeros(B, dtype=torch.int64, device=dev)
    if args.impl == "ours":
        import torch_distlearn_b200 as dl
        comm = dl.P2PCommunicator(rank=rank, world=world, device=dev, heap_mb=16)
        flat = dl.FlatParams(comm, model)
        ea = dl.AllReduceEA(comm, args.tau, args.alpha, flat=flat)
        ea.synchronizeParameters()
        p, gr = flat.params.tensor, flat.grads.tensor
        exchange = ea.averageParameters
    else:
        from baseline.nccl_algos import NcclAllReduceEA, NcclFlat
        flat = NcclFlat(model)
        ea = NcclAllReduceEA(flat, args.tau, args.alpha)
        if world > 1:
            dist.broadcast(flat.params, src=0)
            ea.center.copy_(flat.params)
        p, gr = flat.params, flat.grads
        exchange = ea.average_parameters

    fused = args.impl == "ours" and args.fused and B <= 8
    if fused:
        from torch_distlearn_b200.ops.mnist import mnist_step, model_tensors
        mp, mg = model_tensors(model)
        loss_dev = torch.zeros(1, device=dev)

    def local_step(xb=None, yb=None):
        xb, yb = (x, y) if xb is None else (xb, yb)
        if fused:
            mnist_step(xb, yb, mp, mg, loss_dev, None, lr=args.lr, apply_sgd=True)     # examples/mnist-ea.lua:100-107 in one launch
            return
        F.nll_loss(model(xb), yb).backward()
        local_sgd(p, gr, None, lr=args.lr, zero_grads=True)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            local_step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    G = args.graph_steps or args.tau
    assert args.tau % G == 0 and G <= 16, "--graph-steps must divide tau and be at most 16"
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for k in range(G):
            local_step(xs[k], ys[k])          # every step of the replay trains on its own batch, read in place

    def step(i):
        if i % G == 0:
            graph.replay()
        exchange()                            # counts the step; every tau-th call launches the exchange

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    if rank == 0:
        print(json.dumps({"bench": "mnist_allreduce_ea", "impl": args.impl, "fused_local_step": bool(fused), "mnist_kernel_form": (int(_lib.cuda().dl_mnist_form(0)) if fused else None), "n_gpus": world,
                          "per_gpu_batch": B, "tau": args.tau, "steps_per_graph_replay": G,
                          "alpha": args.alpha, "steps": args.steps, "ms_per_step": round(ms / args.steps, 5),
                          "samples_per_s": round(B * world * args.steps / (ms / 1e3), 1),
                          "exchanges_per_s": round(args.steps / args.tau / (ms / 1e3), 1)}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

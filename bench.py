#!/usr/bin/env python
"""Headline benchmark: CIFAR-10 convnet (examples/Model.lua) AllReduceSGD training throughput.

    python bench.py --gpus N --steps K --warmup W            # this framework (fused sm_100a path)
    python bench.py --impl nccl ...                          # torch.distributed NCCL baseline
    python bench.py --impl reference ...                     # the reference itself (Lua/Torch7: unavailable here)

For N > 1 launch with ``python -m torch.distributed.run --nnodes=1 --nproc-per-node N ...``.
Metric: samples/s over the whole job; synthetic CIFAR-shaped data, random-init weights, fp32 (the
reference casts everything to float, examples/cifar10.lua:136), per-GPU batch fixed (weak
scaling).  ``value`` is device-timed (CUDA events, max over ranks) with the batch already on the
device; ``e2e`` goes through the public TrainStep.step() API and includes, every step, the H2D
copy of that step's batch from pinned host memory and the D2H read of the loss.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--impl", choices=["ours", "nccl", "reference"], default="ours")
    p.add_argument("--per-gpu-batch", type=int, default=128, help="examples/AsyncEASGD.sh:39 uses 128 per GPU")
    p.add_argument("--lr", type=float, default=0.1, help="examples/cifar10.lua:7")
    p.add_argument("--graph", type=int, default=None,
                   help="capture the whole step in a CUDA graph (default: 1 for ours, 0 for nccl - the stock eager path; "
                        "capturing NCCL inside the step graph hung on the 8-GPU box)")
    p.add_argument("--tf32", type=int, default=0, help="allow TF32 convolutions (default: strict fp32 like the reference)")
    p.add_argument("--variant", default="auto")
    p.add_argument("--nblocks", type=int, default=0)
    p.add_argument("--heap-mb", type=float, default=128)
    p.add_argument("--conv", choices=["tc", "cudnn"], default="tc",
                   help="ours: tc = convolutions as implicit GEMMs on tcgen05 (3xTF32, fp32-level accuracy, csrc/tc_gemm.cu); "
                        "cudnn = stock torch modules (what the nccl baseline always uses)")
    p.add_argument("--conv-terms", type=int, default=3, choices=[1, 3], help="3 = 3xTF32 (fp32-accurate), 1 = plain TF32")
    p.add_argument("--flat-order", choices=["module", "execution"], default="module",
                   help="layout of the flat parameter vector; 'execution' lets --bucket-mb buckets leave layer by layer")
    p.add_argument("--bucket-mb", type=float, default=0.0,
                   help="ours: launch the fused kernel per gradient bucket of this size, overlapped with backward (0 = one launch)")
    p.add_argument("--bucket-blocks", type=int, default=32, help="CTAs per bucket launch (leaves the other SMs to cuDNN)")
    a = p.parse_args()
    if a.graph is None:
        a.graph = 1 if a.impl == "ours" else 0
    return a


def reference_arm(args) -> None:
    # The reference is pure Lua/Torch7 (0 Python files, rockspec build type "builtin"); the offline pip
    # install of /root/reference fails for lack of a Python project, and the image has no lua/luajit/th
    # (DESIGN.md "Reference arm").
    if int(os.environ.get("RANK", "0")) == 0:       # one line per job, also under torchrun
        print(json.dumps({"impl": "reference",
                          "unavailable": "reference is Lua/Torch7 + luarocks ipc: not pip-installable, no lua/th runtime in the image"}))


def main() -> None:
    args = parse()
    if args.impl == "reference":
        reference_arm(args)
        return

    import torch
    import torch.distributed as dist
    import torch.nn.functional as F

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (use examples/mnist.py --backend host for the CPU configuration)")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    torch.backends.cudnn.allow_tf32 = bool(args.tf32)
    torch.backends.cuda.matmul.allow_tf32 = bool(args.tf32)
    torch.backends.cudnn.benchmark = True
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if args.impl == "nccl":
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)  # bootstrap only

    from torch_distlearn_b200.models import cifar_convnet, cifar_convnet_tc
    from torch_distlearn_b200.utils.timing import ClockSampler, max_over_ranks

    B = args.per_gpu_batch
    use_tc = args.impl == "ours" and args.conv == "tc"
    model = (cifar_convnet_tc(terms=args.conv_terms) if use_tc else cifar_convnet()).to(dev)
    conv_desc = (f"tcgen05 implicit GEMM, {'3xTF32 (fp32-level accuracy)' if args.conv_terms == 3 else 'TF32'}" if use_tc
                 else f"cuDNN via torch, allow_tf32={bool(args.tf32)}")
    K, W = args.steps, max(args.warmup, 3)
    # a pool of distinct pinned host batches, cycled: every step copies a fresh batch
    pool = 8
    from torch_distlearn_b200.utils.data import cifar_dataset
    ds = cifar_dataset(partition=rank + 1, partitions=world, num_samples=pool * B * world)
    get_batch, _ = ds.sampledBatcher(B, "label-uniform")      # examples/cifar10.lua:53-54
    batches = [get_batch() for _ in range(pool)]
    host_x = [b.input.contiguous().pin_memory() for b in batches]
    host_y = [b.target.contiguous().pin_memory() for b in batches]
    dev_x = [t.to(dev) for t in host_x]
    dev_y = [t.to(dev) for t in host_y]
    h2d_bytes = host_x[0].numel() * 4 + host_y[0].numel() * 8
    d2h_bytes = 4
    launches = 0

    if args.impl == "ours":
        import torch_distlearn_b200 as dl
        from torch_distlearn_b200.parallel.trainer import TrainStep
        comm = dl.P2PCommunicator(rank=rank, world=world, device=dev, heap_mb=args.heap_mb)
        flat = dl.FlatParams(comm, model, order=args.flat_order)
        sgd = dl.AllReduceSGD(comm, flat)
        sgd.synchronizeParameters()
        step = TrainStep(model, flat, sgd, (B, 3, 32, 32), lr=args.lr, use_graph=bool(args.graph), variant=args.variant,
                         nblocks=args.nblocks, bucket_mb=args.bucket_mb, bucket_blocks=args.bucket_blocks)
        step.capture()

        def run_dev(i):
            step.load(dev_x[i % pool], dev_y[i % pool])
            step.run()

        # End to end through the public API: the next batch is prefetched (pinned host -> device on a copy stream)
        # while the current step runs, the loss is read back to the host every step.  Every step's inputs cross
        # PCIe inside the timed region (batch i+1 is copied during step i; one batch is staged before the region).
        def run_e2e(i):
            if not step._staged:
                step.prefetch(host_x[i % pool], host_y[i % pool])
            step.run()
            step.prefetch(host_x[(i + 1) % pool], host_y[(i + 1) % pool])
            return step.loss_value()

        describe = comm.describe()
    else:
        from baseline.nccl_algos import NcclAllReduceSGD, NcclFlat
        flat = NcclFlat(model)
        algo = NcclAllReduceSGD(flat)
        algo.synchronize_parameters()
        x_static = torch.zeros(B, 3, 32, 32, device=dev)
        y_static = torch.zeros(B, dtype=torch.int64, device=dev)
        loss_dev = torch.zeros((), device=dev)
        loss_host = torch.zeros(()).pin_memory()

        def body():
            out = model(x_static)
            loss = F.nll_loss(out, y_static)
            loss.backward()
            loss_dev.copy_(loss.detach())
            algo.step(args.lr)

        nccl_graph = None
        if args.graph:
            # the strong baseline: the same whole-step CUDA graph, with NCCL's all-reduce captured in it
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(3):
                    body()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            nccl_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(nccl_graph):
                body()
            torch.cuda.synchronize(dev)

        def run_body():
            if nccl_graph is not None:
                nccl_graph.replay()
            else:
                body()

        def run_dev(i):
            x_static.copy_(dev_x[i % pool], non_blocking=True)
            y_static.copy_(dev_y[i % pool], non_blocking=True)
            run_body()

        def run_e2e(i):
            x_static.copy_(host_x[i % pool], non_blocking=True)
            y_static.copy_(host_y[i % pool], non_blocking=True)
            run_body()
            loss_host.copy_(loss_dev, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return float(loss_host)

        describe = {"backend": "nccl"}

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-timed region -------------------------------------------------------------------
    for i in range(W):
        run_dev(i)
    barrier()
    sampler = ClockSampler(gpu_index=local_rank, period_ms=100).start() if rank == 0 else None
    if args.impl == "ours":
        l0 = step.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(K):
        run_dev(i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    if args.impl == "ours":
        launches = step.kernel_launches - l0      # every kernel of libdistlearn_b200 in the timed region (graph replays included)
    ms = max_over_ranks(ms, world)
    # ---- end-to-end region (public API, H2D + D2H every step) -------------------------------------
    for i in range(3):
        run_e2e(i)
    barrier()
    t0 = time.perf_counter()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    last = 0.0
    for i in range(K):
        last = run_e2e(i)
    e3.record()
    barrier()
    e2e_ms = max_over_ranks(max(e2.elapsed_time(e3), 0.0), world)
    wall_ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world)
    clocks = sampler.stop() if sampler is not None else None

    if rank == 0:
        total = B * world
        value = total * K / (ms / 1e3)
        e2e_value = total * K / (max(e2e_ms, 1e-9) / 1e3)
        base = None
        try:
            with open(os.path.join(ROOT, "baseline", "measured.json")) as f:
                base = json.load(f).get("cifar_sgd_samples_per_s", {}).get(str(world))
        except Exception:
            base = None
        out = {
            "metric": "CIFAR-10 convnet AllReduceSGD training throughput", "value": round(value, 1), "unit": "samples/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(ms / K, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": (round(value / base, 4) if base else None), "dtype": "fp32",
            "data": "synthetic CIFAR-shaped 3x32x32, random-init weights", "impl": args.impl,
            "config": {"model": "cifar10_convnet (examples/Model.lua, 4,328,970 params)", "global_batch": total,
                       "per_gpu_batch": B, "seq_len": None, "parallelism": f"dp{world}", "lr": args.lr,
                       "tf32": bool(args.tf32), "conv": conv_desc, "cuda_graph": bool(args.graph), "bucket_mb": args.bucket_mb, "flat_order": args.flat_order,
                       "l2": "no explicit flush: each step streams >126 MB of activations/gradients (working set > L2)",
                       "transport": describe},
            "e2e": {"value": round(e2e_value, 1), "unit": "samples/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h_bytes, "ms_per_step": round(e2e_ms / K, 4),
                    "input_pipeline": ("next batch prefetched on a copy stream during the current step (TrainStep.prefetch)"
                                       if args.impl == "ours" else "copy in front of the step"),
                    "wall_ms_per_step": round(wall_ms / K, 4), "last_loss": round(last, 5)},
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

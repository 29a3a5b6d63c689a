#!/usr/bin/env python
"""Headline benchmark: CIFAR-10 convnet (examples/Model.lua) data-parallel training throughput.

    python bench.py --gpus N --steps K --warmup W            # this framework (fused sm_100a path) + same-lease NCCL arms
    python bench.py --impl nccl [--nccl-variant flat|ddp]    # torch.distributed NCCL baseline only
    python bench.py --impl reference ...                     # the reference itself (Lua/Torch7: unavailable here)
    python bench.py --algo ea                                # AllReduceEA (tau 10, alpha 0.2) instead of AllReduceSGD
    python bench.py --global-batch 32                        # the reference's own configuration: batch 32 split over the
                                                             # nodes (examples/cifar10.lua:6,36) = strong scaling

For N > 1 launch with ``python -m torch.distributed.run --nnodes=1 --nproc-per-node N ...``.
Metric: samples/s over the whole job; synthetic CIFAR-shaped data, random-init weights, fp32 (the
reference casts everything to float, examples/cifar10.lua:136).  ``value`` is device-timed (CUDA
events, max over ranks) with the batch already on the device; ``e2e`` goes through the public
TrainStep.step() API and includes, every step, the H2D copy of that step's batch from pinned host
memory and the D2H read of the loss.

Baselines are measured IN THE SAME PROCESS, right after our arm, with the same steps and warm-up and their own clock
record (``baselines``): the stock torch path — cuDNN convolutions, NCCL ``all_reduce(AVG)`` on the flat gradient,
``p.add_`` — once in strict fp32 (what the reference computes in) and once with cuDNN allowed TF32 (torch's default),
plus, for N > 1, DistributedDataParallel (bucketed all-reduce overlapped with backward) as the strong NCCL variant.
``vs_baseline`` divides by the best strict-fp32 arm of this very run; ``vs_baseline_tf32`` by the TF32 arm.
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
    p.add_argument("--algo", choices=["sgd", "ea"], default="sgd",
                   help="sgd: AllReduceSGD (headline); ea: AllReduceEA tau/alpha (examples/client_remote.lua:192-258)")
    p.add_argument("--tau", type=int, default=10)
    p.add_argument("--alpha", type=float, default=0.2)
    p.add_argument("--per-gpu-batch", type=int, default=128, help="examples/AsyncEASGD.sh:39 uses 128 per GPU")
    p.add_argument("--global-batch", type=int, default=0,
                   help="> 0: total batch split over the GPUs, ceil(batch/N) each (examples/cifar10.lua:36) - strong scaling")
    p.add_argument("--lr", type=float, default=0.1, help="examples/cifar10.lua:7")
    p.add_argument("--graph", type=int, default=None,
                   help="capture the whole step in a CUDA graph (default: 1 for ours, 0 for nccl - the stock eager path; "
                        "capturing NCCL inside the step graph hung on the 8-GPU box)")
    p.add_argument("--tf32", type=int, default=0, help="--impl nccl: allow TF32 convolutions (default: strict fp32 like the reference)")
    p.add_argument("--nccl-variant", choices=["flat", "ddp"], default="flat")
    p.add_argument("--baselines", type=int, default=1, help="--impl ours: also measure the NCCL arms in this process")
    p.add_argument("--variant", default="auto")
    p.add_argument("--nblocks", type=int, default=0)
    p.add_argument("--heap-mb", type=float, default=128)
    p.add_argument("--conv", choices=["tc", "cudnn"], default="tc",
                   help="ours: tc = convolutions as implicit GEMMs on tcgen05 (3xTF32, fp32-level accuracy, csrc/tc_gemm.cu); "
                        "cudnn = stock torch modules (what the nccl baseline always uses)")
    p.add_argument("--conv-terms", type=int, default=3, choices=[1, 3], help="3 = 3xTF32 (fp32-accurate), 1 = plain TF32")
    p.add_argument("--flat-order", choices=["module", "execution"], default=None,
                   help="layout of the flat parameter vector; 'execution' lets --bucket-mb buckets leave layer by layer "
                        "(default: execution when buckets are on)")
    p.add_argument("--bucket-mb", type=float, default=None,
                   help="ours: launch the fused kernel per gradient bucket of this size, overlapped with backward "
                        "(0 = one launch after backward; default: 0.5 for N >= 4 = one bucket per layer of this model)")
    p.add_argument("--bucket-blocks", type=int, default=0,
                   help="CTAs per bucket launch (0 = one light CTA per SM, beside the convolution's CTA)")
    p.add_argument("--nvls-permille", type=int, default=-1,
                   help="share of every slice the hybrid / light kernels reduce in the switch (-1: the measured table's value)")
    a = p.parse_args()
    if a.nvls_permille >= 0:
        os.environ["DL_B200_NVLS_PERMILLE"] = str(a.nvls_permille)
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


class Arm:
    """One implementation under test: run_dev(i) (batch already on the device), run_e2e(i) -> loss (host batch in,
    loss out), launches() -> kernels of this library so far."""
    name = ""
    describe: dict = {}

    def launches(self) -> int:
        return 0


def measure(arm: Arm, K: int, W: int, world: int, rank: int, local_rank: int, dev, total_batch: int):
    import torch
    import torch.distributed as dist
    from torch_distlearn_b200.utils.timing import ClockSampler, max_over_ranks

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-timed region -------------------------------------------------------------------
    for i in range(W):
        arm.run_dev(i)
    barrier()
    sampler = ClockSampler(gpu_index=local_rank, period_ms=100).start() if rank == 0 else None
    l0 = arm.launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(K):
        arm.run_dev(i)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    launches = arm.launches() - l0
    # ---- end-to-end region (public API, H2D + D2H every step) -------------------------------------
    for i in range(3):
        arm.run_e2e(i)
    barrier()
    t0 = time.perf_counter()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    last = 0.0
    for i in range(K):
        last = arm.run_e2e(i)
    e3.record()
    barrier()
    e2e_ms = max_over_ranks(max(e2.elapsed_time(e3), 0.0), world)
    wall_ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world)
    clocks = sampler.stop() if sampler is not None else None
    return {"value": total_batch * K / (ms / 1e3), "ms_per_step": ms / K, "e2e_value": total_batch * K / (max(e2e_ms, 1e-9) / 1e3),
            "e2e_ms_per_step": e2e_ms / K, "wall_ms_per_step": wall_ms / K, "last_loss": last, "launches": launches, "clocks": clocks}


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
    torch.backends.cudnn.benchmark = True
    nccl_group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if args.impl == "nccl":
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)  # bootstrap only
    if args.bucket_mb is None:
        # per-layer buckets under backward pay where the light kernel can reduce in the switch (measured on 8 GPUs: 898k
        # samples/s against 879k without buckets); with two GPUs the whole-vector kernel after backward is as fast
        args.bucket_mb = 0.5 if (world >= 4 and args.algo == "sgd") else 0.0
    if args.flat_order is None:
        args.flat_order = "execution" if args.bucket_mb > 0 else "module"

    from torch_distlearn_b200.models import cifar_convnet, cifar_convnet_tc

    strong = args.global_batch > 0
    B = -(-args.global_batch // world) if strong else args.per_gpu_batch      # ceil(batchSize / numNodes)
    total = B * world
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

    # ------------------------------------------------------------------------------------------------ ours
    def make_ours() -> Arm:
        import torch_distlearn_b200 as dl
        from torch_distlearn_b200.parallel.trainer import EATrainStep, TrainStep
        use_tc = args.conv == "tc"
        model = (cifar_convnet_tc(terms=args.conv_terms) if use_tc else cifar_convnet()).to(dev)
        comm = dl.P2PCommunicator(rank=rank, world=world, device=dev, heap_mb=args.heap_mb)
        flat = dl.FlatParams(comm, model, order=args.flat_order)
        if args.algo == "sgd":
            sgd = dl.AllReduceSGD(comm, flat)
            sgd.synchronizeParameters()
            step = TrainStep(model, flat, sgd, (B, 3, 32, 32), lr=args.lr, use_graph=bool(args.graph), variant=args.variant,
                             nblocks=args.nblocks, bucket_mb=args.bucket_mb, bucket_blocks=args.bucket_blocks)
        else:
            ea = dl.AllReduceEA(comm, tau=args.tau, alpha=args.alpha, flat=flat)
            ea.synchronizeParameters()
            step = EATrainStep(model, flat, ea, (B, 3, 32, 32), lr=args.lr, use_graph=bool(args.graph))
        step.load(dev_x[0], dev_y[0])       # warm up on a real batch, not on zeros
        step.capture()
        arm = Arm()
        arm.name = "ours"
        arm.conv_desc = (f"tcgen05 implicit GEMM, {'3xTF32 (fp32-level accuracy)' if args.conv_terms == 3 else 'TF32'}" if use_tc
                         else "cuDNN via torch, allow_tf32=False")
        arm.describe = comm.describe()
        arm.describe["variant_table"] = comm._variants.source

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

        arm.run_dev, arm.run_e2e = run_dev, run_e2e
        arm.launches = lambda: step.kernel_launches      # every kernel of libdistlearn_b200 (graph replays included)
        arm.pipeline = "next batch prefetched on a copy stream during the current step (TrainStep.prefetch)"
        arm.step = step
        return arm

    # ------------------------------------------------------------------------------------------------ NCCL arms
    def make_nccl(tf32: bool, variant: str, group) -> Arm:
        from baseline.nccl_algos import NcclAllReduceEA, NcclAllReduceSGD, NcclFlat
        torch.backends.cudnn.allow_tf32 = bool(tf32)
        torch.backends.cuda.matmul.allow_tf32 = bool(tf32)
        model = cifar_convnet().to(dev)
        x_static = torch.zeros(B, 3, 32, 32, device=dev)
        y_static = torch.zeros(B, dtype=torch.int64, device=dev)
        loss_dev = torch.zeros((), device=dev)
        loss_host = torch.zeros(()).pin_memory()
        arm = Arm()
        arm.name = f"nccl_{variant}_{'tf32' if tf32 else 'fp32'}"
        arm.conv_desc = f"cuDNN via torch, allow_tf32={bool(tf32)}"
        arm.describe = {"backend": "nccl", "variant": variant}
        if variant == "ddp" and world > 1:
            # the strong library baseline: bucketed all-reduce (25 MB buckets by default, the first one smaller) launched
            # from autograd hooks and overlapped with the rest of backward, then a fused multi-tensor SGD
            from torch.nn.parallel import DistributedDataParallel as DDP
            ddp = DDP(model, device_ids=[dev.index], process_group=group, gradient_as_bucket_view=True, bucket_cap_mb=4)
            opt = torch.optim.SGD(model.parameters(), lr=args.lr, foreach=True)

            def body():
                out = ddp(x_static)
                loss = F.nll_loss(out, y_static)
                loss.backward()
                loss_dev.copy_(loss.detach())
                opt.step()
                opt.zero_grad(set_to_none=False)
        else:
            flat = NcclFlat(model)
            if args.algo == "sgd":
                algo = NcclAllReduceSGD(flat, group)
                algo.synchronize_parameters()
            else:
                algo = NcclAllReduceEA(flat, args.tau, args.alpha, group)

            def body():
                out = model(x_static)
                loss = F.nll_loss(out, y_static)
                loss.backward()
                loss_dev.copy_(loss.detach())
                if args.algo == "sgd":
                    algo.step(args.lr)
                else:
                    flat.params.add_(flat.grads, alpha=-args.lr)       # local SGD, then the elastic average every tau steps
                    flat.grads.zero_()
                    algo.average_parameters()

        nccl_graph = None
        if args.graph and args.impl == "nccl":
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

        arm.run_dev, arm.run_e2e = run_dev, run_e2e
        arm.pipeline = "copy in front of the step"
        return arm

    def nccl_pg():
        nonlocal nccl_group
        if world > 1 and nccl_group is None:
            nccl_group = dist.group.WORLD if args.impl == "nccl" else dist.new_group(backend="nccl")
        return nccl_group

    if args.impl == "ours":
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        arm = make_ours()
    else:
        arm = make_nccl(bool(args.tf32), args.nccl_variant, nccl_pg())
    res = measure(arm, K, W, world, rank, local_rank, dev, total)

    baselines = {}
    if args.impl == "ours" and args.baselines:
        plans = [("nccl_fp32", False, "flat"), ("nccl_tf32", True, "flat")]
        if world > 1:
            plans.append(("nccl_ddp_fp32", False, "ddp"))
        for key, tf32, variant in plans:
            try:
                b = measure(make_nccl(tf32, variant, nccl_pg()), K, W, world, rank, local_rank, dev, total)
                baselines[key] = {"value": round(b["value"], 1), "ms_per_step": round(b["ms_per_step"], 4),
                                  "e2e_value": round(b["e2e_value"], 1), "clocks": b["clocks"],
                                  "what": ("DistributedDataParallel (bucketed all-reduce under backward) + foreach SGD" if variant == "ddp"
                                           else "all_reduce(AVG) of the flat gradient + p.add_") + f", cuDNN allow_tf32={tf32}, eager"}
            except Exception as e:  # noqa: BLE001 - a failing baseline must not take the headline down with it
                baselines[key] = {"error": repr(e)[:200]}
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False

    if rank == 0:
        fp32_arms = [v["value"] for k, v in baselines.items() if k.endswith("fp32") and "value" in v]
        base_fp32 = max(fp32_arms) if fp32_arms else None
        base_tf32 = baselines.get("nccl_tf32", {}).get("value")
        value = res["value"]
        metric = ("CIFAR-10 convnet AllReduceSGD training throughput" if args.algo == "sgd"
                  else f"CIFAR-10 convnet AllReduceEA (tau {args.tau}, alpha {args.alpha}) training throughput")
        out = {
            "metric": metric, "value": round(value, 1), "unit": "samples/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(res["ms_per_step"], 4), "higher_is_better": True,
            "scaling": "strong" if strong else "weak",
            "vs_baseline": (round(value / base_fp32, 4) if base_fp32 else None),
            "vs_baseline_fp32": (round(value / base_fp32, 4) if base_fp32 else None),
            "vs_baseline_tf32": (round(value / base_tf32, 4) if base_tf32 else None),
            "baseline_note": ("the reference publishes no number (BASELINE.md §1) and cannot run here; vs_baseline divides by the best "
                              "strict-fp32 torch+NCCL arm measured in this same process and lease (see `baselines`)"),
            "dtype": "fp32",
            "data": "synthetic CIFAR-shaped 3x32x32, random-init weights", "impl": args.impl,
            "config": {"model": "cifar10_convnet (examples/Model.lua, 4,328,970 params)", "global_batch": total,
                       "per_gpu_batch": B, "seq_len": None, "parallelism": f"dp{world}", "lr": args.lr, "algo": args.algo,
                       "tf32": bool(args.tf32) if args.impl == "nccl" else False, "conv": arm.conv_desc, "cuda_graph": bool(args.graph),
                       "bucket_mb": args.bucket_mb, "flat_order": args.flat_order,
                       "l2": "no explicit flush: each step streams >126 MB of activations/gradients (working set > L2)"
                             if B >= 64 else "no explicit flush; at this batch size the working set fits the L2 (latency-bound regime)",
                       "transport": arm.describe},
            "e2e": {"value": round(res["e2e_value"], 1), "unit": "samples/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h_bytes, "ms_per_step": round(res["e2e_ms_per_step"], 4),
                    "input_pipeline": arm.pipeline,
                    "wall_ms_per_step": round(res["wall_ms_per_step"], 4), "last_loss": round(res["last_loss"], 5)},
            "gpu_launches": int(res["launches"]),
            "clocks": res["clocks"],
            "baselines": baselines,
        }
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Measured NVLink / NVSwitch peaks of THIS box, the denominators of every roofline fraction in
profiles/sweep_*.md.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29561 \
        bench/peak.py [--mb 256] [--out gpurun_out/nvlink_peak_N.json]

Patterns (per-GPU, per-direction GB/s; time = CUDA events around K back-to-back launches, max over ranks):
  uni pull / push      only rank 0 moves data (from / to rank 1): one direction of one GPU's links loaded
  ring pull / push     every rank moves S from / to its right neighbour at once: every GPU sends S and receives S
  all pull             every rank reads S/(N-1) from each peer: the traffic shape of the two-shot kernels
  copy engine          cudaMemcpyAsync between heaps (ring)
  multicast store      every rank multimem.st's S/N (all-gather through the switch): ingress S per GPU
  multicast reduce     every rank multimem.ld_reduce's S/N (reduce-scatter in the switch): egress S per GPU
The value written as ``p2p_copy_gbs_per_direction`` — what bench/sweep.py divides by — is the ring pull figure (kernel
loads, both directions loaded), the access pattern of the collectives themselves.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=float, default=256)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import torch_distlearn_b200 as dl
    from torch_distlearn_b200 import _lib
    n = int(args.mb * (1 << 20)) // 4 // (8 * world) * (8 * world)
    comm = dl.P2PCommunicator(rank=rank, world=world, device=dev, heap_mb=args.mb * 2 + 64)
    a = comm.alloc(n, torch.float32, "peak.a")
    b = comm.alloc(n, torch.float32, "peak.b")
    a.tensor.normal_()
    lib, cref = _lib.cuda(), C.byref(comm._c)
    stream = torch.cuda.current_stream(dev).cuda_stream
    nbytes = n * 4
    right = (rank + 1) % world

    def sync():
        torch.cuda.synchronize()
        dist.barrier()

    def timed(fn, participates=True):
        for _ in range(3):
            if participates:
                fn()
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if participates:
            for _ in range(args.iters):
                fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters if participates else 0.0
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sync()
        return float(t.item()) * 1e-3       # seconds per launch, max over ranks

    def stream_fn(mode, peer, nfl, src_off, dst_off):
        return lambda: _lib.check(lib.dl_peak_stream(cref, mode, peer, src_off, dst_off, nfl, 0, stream))

    res = {"world": world, "bytes": nbytes, "transport": comm.describe()}
    # one direction of one GPU
    for name, mode in (("uni_pull", 0), ("uni_push", 1)):
        t = timed(stream_fn(mode, 1 % world, n, a.offset, b.offset), participates=(rank == 0))
        res[name + "_gbs"] = round(nbytes / t / 1e9, 1)
    # every GPU sends S and receives S
    for name, mode in (("ring_pull", 0), ("ring_push", 1)):
        t = timed(stream_fn(mode, right, n, a.offset, b.offset))
        res[name + "_gbs"] = round(nbytes / t / 1e9, 1)
    # all-to-all pull: S/(N-1) from each peer
    if world > 2:
        part = n // (world - 1) // 8 * 8

        def all_pull():
            k = 0
            for p in range(world):
                if p != rank:
                    _lib.check(lib.dl_peak_stream(cref, 0, p, a.offset + k * part * 4, b.offset + k * part * 4, part, 0, stream))
                    k += 1
        t = timed(all_pull)
        res["all_pull_gbs"] = round(part * (world - 1) * 4 / t / 1e9, 1)
    for name, pull in (("copy_engine_pull", 1), ("copy_engine_push", 0)):
        t = timed(lambda pull=pull: _lib.check(lib.dl_peak_memcpy(cref, right, pull, a.offset, b.offset, nbytes, stream)))
        res[name + "_gbs"] = round(nbytes / t / 1e9, 1)
    if comm.has_multicast:
        share = n // world
        off = rank * share * 4
        t = timed(stream_fn(2, rank, share, a.offset + off, b.offset + off))
        res["mc_store_ingress_gbs"] = round(nbytes / t / 1e9, 1)          # every GPU receives S
        res["mc_store_egress_gbs"] = round(share * 4 / t / 1e9, 1)
        t = timed(stream_fn(3, rank, share, a.offset + off, b.offset + off))
        res["mc_reduce_egress_gbs"] = round(nbytes / t / 1e9, 1)          # every GPU is read S (its share N times over)
        res["mc_reduce_result_gbs"] = round(share * 4 / t / 1e9, 1)
        # a single rank multicasting alone (the shape of a root-only NVLS broadcast)
        t = timed(stream_fn(2, rank, n, a.offset, b.offset), participates=(rank == 0))
        res["mc_store_single_root_gbs"] = round(nbytes / t / 1e9, 1)
    res["p2p_copy_gbs_per_direction"] = res["ring_pull_gbs"]
    if rank == 0:
        print(json.dumps(res, indent=1), flush=True)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            json.dump(res, open(args.out, "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

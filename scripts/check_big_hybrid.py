#!/usr/bin/env python
"""Value check of the switch share of the hybrid kernels on both sides of DL_PIPE_MC_BIG (csrc/dl_pipe.cuh), real GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/check_big_hybrid.py

Inputs are small integers, so every sum is exact in fp32 and the results must match bit for bit."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("gloo")
import torch_distlearn_b200 as dl

comm = dl.P2PCommunicator(heap_mb=640)
dev = comm.device
ok = True
for n in ((1 << 20) + 64, (1 << 24) + 64):
    for permille in (1000, 600):
        if not comm.has_multicast:
            continue
        comm._variants.hybrid_permille = {op: permille for op in comm._variants.OPS}
        mark = comm._cursor
        g = comm.alloc(n, torch.float32, "g")
        p = comm.alloc(n, torch.float32, "p")
        idx = torch.arange(n, device=dev, dtype=torch.float32) % 64
        g.tensor.copy_((idx + rank) * world)              # sum over ranks / world = world*idx + sum(r) : exact
        p.tensor.copy_(idx)
        comm.fused_sgd(g, p, None, n, lr=0.5, variant="hybrid")
        torch.cuda.synchronize()
        want = idx - 0.5 * (world * idx + world * (world - 1) / 2)
        good = bool(torch.equal(p.tensor, want)) and float(g.tensor.abs().max()) == 0.0
        b = comm.alloc(n, torch.float32, "b")
        b.tensor.copy_(idx + rank)
        comm.allreduce(b, n, variant="hybrid")
        torch.cuda.synchronize()
        good = good and bool(torch.equal(b.tensor, world * idx + world * (world - 1) / 2))
        flags = [None] * world
        dist.all_gather_object(flags, good)
        if rank == 0:
            print(f"n={n} permille={permille}: {'ok' if all(flags) else 'MISMATCH ' + str(flags)}", flush=True)
        ok = ok and all(flags)
        comm._cursor = mark
if rank == 0:
    print("BIG HYBRID OK" if ok else "BIG HYBRID FAILED", flush=True)
comm.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)

#!/bin/bash
# Round 2, last GPU minutes: the push path (LL) with the contiguous packet layout, 4 GPUs, graph-timed, our kernels only.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 100 $TR --master-port 29542 bench/sweep.py --max-mb 0.0625 --variants ll,twoshot --no-nccl --graph --out gpurun_out/r2x_sweep_4_ll_graph.json > gpurun_out/r2x_sweep_4_ll_graph.log 2>&1; echo "sweep rc=$?"
grep '"impl": "ll"' gpurun_out/r2x_sweep_4_ll_graph.log | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['op'], r['bytes'], r['us'])"

#!/bin/bash
# Round 2, last 8-GPU call: value check and graph-timed sweep of the hybrid kernels with the size-dependent number of
# switch loads in flight (csrc/dl_pipe.cuh), and the default bench with the merged world-8 table.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 120 $TR --master-port 29571 scripts/check_big_hybrid.py > gpurun_out/r2y_check_big_hybrid.log 2>&1; echo "check rc=$?"; grep -E "ok|MISMATCH|OK|FAILED|Error" gpurun_out/r2y_check_big_hybrid.log | head -8
timeout 200 $TR --master-port 29542 bench/sweep.py --ops fused_sgd,allreduce,fused_ea --max-mb 64 --permille 750,1000 --graph --no-nccl --out gpurun_out/r2y_sweep_8_graph.json > gpurun_out/r2y_sweep_8_graph.log 2>&1; echo "sweep rc=$?"
grep -E "hybrid1000|hybrid750" gpurun_out/r2y_sweep_8_graph.log | grep -E "67108864|16777216|1048576 " | cut -c1-200 | head -12

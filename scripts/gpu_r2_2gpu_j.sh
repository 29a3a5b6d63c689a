#!/bin/bash
# Round 2, 2-GPU call j: do per-layer buckets pay at 2 GPUs when the light kernel reduces through the switch?
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
b() { name=$1; shift; timeout 200 $TR --master-port 29551 bench.py --gpus 2 --steps 100 --warmup 10 --baselines 0 "$@" > gpurun_out/r2j_bench2_$name.log 2>&1; echo "bench $name rc=$? $(tail -n 1 gpurun_out/r2j_bench2_$name.log | cut -c1-160)"; }
b b0 --bucket-mb 0
DL_B200_LIGHT_PERMILLE=1000 b exe05_p1000 --bucket-mb 0.5 --flat-order execution
DL_B200_LIGHT_PERMILLE=0 b exe05_p0 --bucket-mb 0.5 --flat-order execution
DL_B200_NVLS_PERMILLE=1000 DL_B200_VARIANT=hybrid b b0_hybrid1000 --bucket-mb 0

#!/bin/bash
# One process per GPU, node i on GPU i (reference: examples/cifar10-cuda.sh pins 4 GPUs; default here: all 8)
cd "$(dirname "$0")"
n=${NUM_GPUS:-8}
export DISTLEARN_JOB=${DISTLEARN_JOB:-$$}
for i in $(seq 1 $n); do
  python cifar10.py --numNodes $n --nodeIndex $i --cuda --gpu $i --batchSize $((128 * n)) --graph "$@" &
done
wait

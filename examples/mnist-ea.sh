#!/bin/bash
# 4 CPU processes, elastic averaging tau=10 alpha=0.2 (reference: examples/mnist-ea.sh)
cd "$(dirname "$0")"
export DISTLEARN_JOB=${DISTLEARN_JOB:-$$}
for i in 1 2 3 4; do
  python mnist_ea.py --numNodes 4 --nodeIndex $i "$@" &
done
wait

#!/bin/bash
# 4 CPU processes (reference: examples/cifar10.sh)
cd "$(dirname "$0")"
export DISTLEARN_JOB=${DISTLEARN_JOB:-$$}
for i in 1 2 3 4; do
  python cifar10.py --numNodes 4 --nodeIndex $i "$@" &
done
wait

#!/bin/bash
# 4 CPU processes on this machine (reference: examples/mnist.sh)
cd "$(dirname "$0")"
export DISTLEARN_JOB=${DISTLEARN_JOB:-$$}
for i in 1 2 3 4; do
  python mnist.py --numNodes 4 --nodeIndex $i "$@" &
done
wait

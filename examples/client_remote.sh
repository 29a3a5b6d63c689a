#!/bin/bash
# Two nodes that need not share a machine or a device type (reference: examples/client_remote.sh runs
# one CPU node with OMP_NUM_THREADS=4 and one GPU node).  Start node 1 on the rendezvous host and
# node 2 anywhere that can reach it:   HOST=10.0.0.1 ./client_remote.sh
cd "$(dirname "$0")"
host=${HOST:-127.0.0.1}
port=${PORT:-8080}
OMP_NUM_THREADS=4 python client_remote.py --backend dist --numNodes 2 --nodeIndex 1 --host $host --port $port "$@" &
if python -c "import torch,sys; sys.exit(0 if torch.cuda.is_available() else 1)"; then
  python client_remote.py --backend dist --numNodes 2 --nodeIndex 2 --host $host --port $port --cuda --gpu 1 "$@" &
else
  OMP_NUM_THREADS=4 python client_remote.py --backend dist --numNodes 2 --nodeIndex 2 --host $host --port $port "$@" &
fi
wait

#!/bin/bash
# Asynchronous EASGD: one center server, one tester and N clients (counterpart of
# examples/AsyncEASGD.sh in the reference: server + tester + 2 GPU clients on one host).
#   ./AsyncEASGD.sh [port] [numClients] [--cuda]
# With --cuda the server's center lives in GPU 1's memory, the tester runs on GPU 2 and client i on
# GPU 2+i; without it everything runs on CPU over shared memory.
cd "$(dirname "$0")"
port=${1:-8080}
numNodes=${2:-2}
cuda=${3:-}
common="--numNodes $numNodes --port $port --tester --communicationTime ${TAU:-10} --testTime ${TESTTIME:-20} --numEpochs ${EPOCHS:-2} \
  --trainSamples ${TRAIN:-2048} --testSamples ${TEST:-512} --batchSize ${BATCH:-128} --save ${SAVE:-Results/easgd}"
export DISTLEARN_JOB=${DISTLEARN_JOB:-$$}
gpu() { if [ -n "$cuda" ]; then echo "--cuda --gpu $1"; fi; }

python easgd_server.py $common --server $(gpu 1) &
pids="$!"
python easgd_tester.py $common $(gpu 2) &
pids="$pids $!"
for i in $(seq 1 $numNodes); do
  python easgd_client.py $common --nodeIndex $i $(gpu $((i + 2))) &
  pids="$pids $!"
done
# clients on other hosts of the same NVLink domain would be started the same way over ssh
rc=0
for p in $pids; do wait $p || rc=$?; done
exit $rc

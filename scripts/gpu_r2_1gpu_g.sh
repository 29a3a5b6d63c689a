#!/bin/bash
# Round 2, 1-GPU call g: one CPU node + one GPU node in one job (examples/client_remote.sh), then ONE ncu capture of the
# MNIST kernel and the tail kernels.
mkdir -p gpurun_out
( cd examples && PORT=8391 timeout 300 bash client_remote.sh --numEpochs 1 --trainSamples 512 --testSamples 128 --batchSize 32 --learningRate 0.01 --tau 2 ) > gpurun_out/r2g_client_remote_mixed.log 2>&1
echo "client_remote.sh (CPU node + GPU node) rc=$?"; tail -n 4 gpurun_out/r2g_client_remote_mixed.log | cut -c1-200
timeout 300 python scripts/ncu_tails_target.py > gpurun_out/r2g_tails_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"mnist_step|bn_relu_pool_bwd|bn_bwd_finalize|colsum|head_|im2col" \
    -c 30 -o gpurun_out/prof_tails -f python scripts/ncu_tails_target.py > gpurun_out/r2g_tails_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2g_tails_plain.log | cut -c1-120; tail -n 2 gpurun_out/r2g_tails_ncu.log | cut -c1-200

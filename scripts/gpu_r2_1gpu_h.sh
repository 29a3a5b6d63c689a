#!/bin/bash
# Round 2, 1-GPU call h: the second form of the MNIST kernel (tests on both forms, bench on both, one ncu capture), and the
# CPU node + GPU node run of examples/client_remote.sh with every node printing its device.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_mnist_fused_gpu.py -m gpu -x -q > gpurun_out/r2h_mnist_tests.log 2>&1; echo "mnist tests rc=$? $(tail -n 1 gpurun_out/r2h_mnist_tests.log)"
for f in 2 1; do
DL_B200_MNIST_FORM=$f timeout 120 python bench/bench_ea.py --impl ours --batch 1 --steps 2000 > gpurun_out/r2h_mnist_ea1_form$f.log 2>&1; echo "mnist ea form $f rc=$? $(tail -n 1 gpurun_out/r2h_mnist_ea1_form$f.log | cut -c1-300)"
done
( cd examples && PORT=8391 timeout 300 bash client_remote.sh --numEpochs 1 --trainSamples 512 --testSamples 128 --batchSize 32 --learningRate 0.01 --tau 2 ) > gpurun_out/r2h_client_remote_mixed.log 2>&1
echo "client_remote.sh (CPU node + GPU node) rc=$?"; tail -n 6 gpurun_out/r2h_client_remote_mixed.log | cut -c1-200
timeout 200 python scripts/ncu_tails_target.py > gpurun_out/r2h_tails_plain.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"mnist_step" -c 8 -o gpurun_out/prof_mnist2 -f python scripts/ncu_tails_target.py > gpurun_out/r2h_mnist_ncu.log 2>&1
echo "ncu rc=$?"; tail -n 2 gpurun_out/r2h_mnist_ncu.log | cut -c1-200

#!/bin/bash
# Round 2, 1-GPU validation call: every GPU test (no -x: all failures in one call), the stream-K decision, the bench.
#   gpurun --timeout 1200 -- 'bash scripts/gpu_r2_1gpu.sh'
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2_all.log 2>&1; echo "pytest rc=$?"; tail -n 25 gpurun_out/r2_all.log | cut -c1-240
DL_B200_TEST_STREAMK=1 timeout 300 python -m pytest tests/test_tc_streamk_gpu.py -m gpu -x -q > gpurun_out/r2_streamk.log 2>&1
echo "streamk rc=$?"; tail -n 3 gpurun_out/r2_streamk.log | cut -c1-240
timeout 200 python bench.py --steps 100 --warmup 10 > gpurun_out/r2_bench1.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/r2_bench1.log | cut -c1-600
DL_B200_TC_STREAMK=1 timeout 200 python bench.py --steps 100 --warmup 10 --baselines 0 > gpurun_out/r2_bench1_streamk.log 2>&1; echo "bench streamk rc=$?"; tail -n 1 gpurun_out/r2_bench1_streamk.log | cut -c1-300
timeout 120 python examples/cifar10.py --cuda --numEpochs 1 --batchSize 128 --trainSamples 8192 --testSamples 1024 --graph > gpurun_out/r2_example.log 2>&1
echo "example rc=$?"; tail -n 3 gpurun_out/r2_example.log | cut -c1-200

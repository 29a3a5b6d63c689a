#!/bin/bash
# Round 2, 1-GPU call k: weights handed to the convolution GEMMs already split (DL_B200_TC_PRESPLIT): tests, then the
# flagship bench and the warm per-group times with it on and off.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_presplit_gpu.py tests/test_tc_gemm_gpu.py tests/test_tc_streamk_gpu.py tests/test_train_gpu.py -m gpu -x -q > gpurun_out/r2k_tests.log 2>&1; echo "tests rc=$? $(tail -n 1 gpurun_out/r2k_tests.log)"; grep -E "^FAILED|^ERROR|Error" gpurun_out/r2k_tests.log | head -8 | cut -c1-300
for p in 1 0; do
DL_B200_TC_PRESPLIT=$p timeout 200 python bench.py --steps 100 --warmup 10 --baselines 0 > gpurun_out/r2k_bench1_presplit$p.log 2>&1; echo "bench presplit=$p rc=$? $(tail -n 1 gpurun_out/r2k_bench1_presplit$p.log | cut -c1-200)"
done
DL_B200_TC_PRESPLIT=1 timeout 200 python scripts/tail_probe.py --out gpurun_out/r2k_tail_probe.json > gpurun_out/r2k_tail_probe.log 2>&1; echo "tail probe rc=$?"; grep "conv forward" gpurun_out/r2k_tail_probe.log | cut -c1-120

#!/bin/bash
# Round 2, 1-GPU call e: warp-wide wgrad producer (numerics + per-GEMM times), per-GEMM times with the deterministic reduction.
mkdir -p gpurun_out
python -m pytest tests/test_tc_gemm_gpu.py tests/test_tc_streamk_gpu.py tests/test_train_gpu.py -m gpu -q -p no:cacheprovider > gpurun_out/r2e_gemm_tests.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r2e_gemm_tests.log | cut -c1-240
timeout 400 python scripts/tail_probe.py --out gpurun_out/tail_probe_e.json > gpurun_out/r2e_tail_probe.log 2>&1; echo "tail probe rc=$?"; grep -E "gemm|conv forward" gpurun_out/r2e_tail_probe.log | cut -c1-120
DL_B200_TC_DETERMINISTIC=1 timeout 400 python scripts/tail_probe.py > gpurun_out/r2e_tail_probe_det.log 2>&1; echo "tail probe det rc=$?"; grep -E "gemm.*terms=3|conv forward" gpurun_out/r2e_tail_probe_det.log | cut -c1-120
timeout 300 python bench.py --steps 100 --warmup 10 --baselines 0 > gpurun_out/r2e_bench1.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/r2e_bench1.log | cut -c1-200
timeout 120 python bench/bench_ea.py --impl ours --steps 2000 > gpurun_out/r2e_mnist_ea1.log 2>&1; echo "mnist ea fused rc=$? $(tail -n 1 gpurun_out/r2e_mnist_ea1.log | cut -c1-300)"
timeout 120 python bench/bench_ea.py --impl nccl --steps 2000 > gpurun_out/r2e_mnist_ea1_nccl.log 2>&1; echo "mnist ea torch rc=$? $(tail -n 1 gpurun_out/r2e_mnist_ea1_nccl.log | cut -c1-300)"

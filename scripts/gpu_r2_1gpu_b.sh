#!/bin/bash
# Round 2, second 1-GPU call: GPU tests, warm per-kernel times of the step, bench, then ONE ncu capture of the kernel
# families that had none (only after the same program exited 0 without ncu).
#   gpurun --timeout 1500 -- 'bash scripts/gpu_r2_1gpu_b.sh'
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2b_all.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/r2b_all.log | cut -c1-240
timeout 300 python scripts/tail_probe.py --out gpurun_out/tail_probe.json > gpurun_out/r2b_tail_probe.log 2>&1; echo "tail probe rc=$?"; cat gpurun_out/r2b_tail_probe.log | cut -c1-120
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/r2b_bench1.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/r2b_bench1.log | cut -c1-300
timeout 300 python bench.py --steps 100 --warmup 10 --algo ea --baselines 0 > gpurun_out/r2b_bench1_ea.log 2>&1; echo "bench ea rc=$?"; tail -n 1 gpurun_out/r2b_bench1_ea.log | cut -c1-300
timeout 300 python scripts/ncu_kernels_target.py > gpurun_out/r2b_ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:"pipe_kernel|ll_kernel|broadcast_wide|allreduce_twoshot|async_ea_client|bn_|colsum|head_|im2col" \
    -c 40 -o gpurun_out/prof_kernels -f python scripts/ncu_kernels_target.py > gpurun_out/r2b_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2b_ncu_plain.log | cut -c1-120; tail -n 3 gpurun_out/r2b_ncu.log | cut -c1-200

#!/bin/bash
# Round 2, fourth 1-GPU call: GPU tests (deterministic stream-K reduction, batched LL polling), bench with and without the
# deterministic reduction, then ONE ncu capture of the three GEMMs of a layer.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2d_all.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/r2d_all.log | cut -c1-240
timeout 300 python bench.py --steps 100 --warmup 10 --baselines 0 > gpurun_out/r2d_bench1.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/r2d_bench1.log | cut -c1-200
DL_B200_TC_DETERMINISTIC=1 timeout 300 python bench.py --steps 100 --warmup 10 --baselines 0 > gpurun_out/r2d_bench1_det.log 2>&1; echo "bench det rc=$?"; tail -n 1 gpurun_out/r2d_bench1_det.log | cut -c1-200
timeout 300 python scripts/ncu_gemm_modes_target.py > gpurun_out/r2d_gemm_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 9 -c 3 -o gpurun_out/prof_gemm_modes -f python scripts/ncu_gemm_modes_target.py > gpurun_out/r2d_gemm_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2d_gemm_plain.log | tail -2; tail -n 2 gpurun_out/r2d_gemm_ncu.log | cut -c1-200

#!/bin/bash
# Round 2, third 1-GPU call: GPU tests, warm per-kernel times, bench.   gpurun --timeout 1200 -- 'bash scripts/gpu_r2_1gpu_c.sh'
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2c_all.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/r2c_all.log | cut -c1-240
timeout 400 python scripts/tail_probe.py --out gpurun_out/tail_probe_c.json > gpurun_out/r2c_tail_probe.log 2>&1; echo "tail probe rc=$?"; cat gpurun_out/r2c_tail_probe.log | cut -c1-120
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/r2c_bench1.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/r2c_bench1.log | cut -c1-300
timeout 120 python bench/bench_ea.py --impl ours --steps 2000 > gpurun_out/r2c_mnist_ea1.log 2>&1; echo "mnist ea fused rc=$? $(tail -n 1 gpurun_out/r2c_mnist_ea1.log | cut -c1-300)"
timeout 120 python bench/bench_ea.py --impl ours --fused 0 --steps 2000 > gpurun_out/r2c_mnist_ea1_unfused.log 2>&1; echo "mnist ea unfused rc=$? $(tail -n 1 gpurun_out/r2c_mnist_ea1_unfused.log | cut -c1-300)"

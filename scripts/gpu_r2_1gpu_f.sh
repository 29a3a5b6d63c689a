#!/bin/bash
# Round 2, 1-GPU call f: full GPU tests after the GEMM revert and the MNIST kernel's parallel reductions; benches.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2f_all.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r2f_all.log | cut -c1-240
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/r2f_bench1.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/r2f_bench1.log | cut -c1-200
timeout 120 python bench/bench_ea.py --impl ours --steps 2000 > gpurun_out/r2f_mnist_ea1.log 2>&1; echo "mnist ea fused rc=$? $(tail -n 1 gpurun_out/r2f_mnist_ea1.log | cut -c1-300)"
timeout 120 python examples/mnist_ea.py --cuda --numEpochs 1 --trainSamples 2048 > gpurun_out/r2f_mnist_ea_example.log 2>&1; echo "mnist_ea example rc=$? $(tail -n 1 gpurun_out/r2f_mnist_ea_example.log | cut -c1-200)"
timeout 120 python examples/mnist_ea.py --cuda --eager --numEpochs 1 --trainSamples 2048 > gpurun_out/r2f_mnist_ea_example_eager.log 2>&1; echo "mnist_ea example eager rc=$? $(tail -n 1 gpurun_out/r2f_mnist_ea_example_eager.log | cut -c1-200)"

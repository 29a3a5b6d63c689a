#!/bin/bash
# Round 2, 1-GPU call i: the switch-reduction kernels on ONE GPU (single-device multicast object), the pull form of the
# MNIST kernel's exchanges.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nvls_one_gpu.py -m gpu -q -rs > gpurun_out/r2i_nvls1_tests.log 2>&1; echo "nvls one-GPU tests rc=$? $(tail -n 1 gpurun_out/r2i_nvls1_tests.log)"; grep -i "skip\|Error" gpurun_out/r2i_nvls1_tests.log | head -5 | cut -c1-300
timeout 300 python -m pytest tests/test_mnist_fused_gpu.py -m gpu -x -q > gpurun_out/r2i_mnist_tests.log 2>&1; echo "mnist tests rc=$? $(tail -n 1 gpurun_out/r2i_mnist_tests.log)"
for f in 2 1; do
DL_B200_MNIST_FORM=$f timeout 120 python bench/bench_ea.py --impl ours --batch 1 --steps 2000 > gpurun_out/r2i_mnist_ea1_form$f.log 2>&1; echo "mnist ea form $f rc=$? $(tail -n 1 gpurun_out/r2i_mnist_ea1_form$f.log | cut -c1-300)"
done
timeout 200 python scripts/ncu_tails_target.py > gpurun_out/r2i_tails_plain.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"mnist_step" -c 8 -o gpurun_out/prof_mnist3 -f python scripts/ncu_tails_target.py > gpurun_out/r2i_mnist_ncu.log 2>&1
echo "ncu rc=$?"

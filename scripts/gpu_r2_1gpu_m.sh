#!/bin/bash
# Round 2, last call: examples/mnist.py on a GPU with the one-launch forward+backward kernel and with torch (--eager).
mkdir -p gpurun_out
cd examples
for mode in "" "--eager"; do
timeout 120 python mnist.py --numNodes 1 --nodeIndex 1 --cuda --numEpochs 1 --trainSamples 4000 --batchSize 1 --reportEvery 100000 $mode > ../gpurun_out/r2m_mnist_example${mode/--/_}.log 2>&1; echo "mnist.py $mode rc=$? $(tail -n 1 ../gpurun_out/r2m_mnist_example${mode/--/_}.log | cut -c1-200)"
done

#!/bin/bash
# First GPU call of the next round (1 GPU, ~3 minutes): everything that was written after this round's GPU minutes
# were spent.  Usage:  gpurun --timeout 400 -- 'bash scripts/gpu_next_round.sh'
# (2-GPU part: gpurun --gpus 2 --timeout 400 -- 'bash scripts/gpu_next_round.sh 2')
mkdir -p gpurun_out
N=${1:-1}
if [ "$N" = "1" ]; then
  python -m pytest tests -m gpu -x -q > gpurun_out/n_all.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/n_all.log
  # experimental stream-K GEMM: numerics first (own process: a trap poisons the context), then the step time with it
  DL_B200_TEST_STREAMK=1 timeout 300 python -m pytest tests/test_tc_streamk_gpu.py -m gpu -x -q > gpurun_out/n_streamk.log 2>&1
  echo "streamk rc=$?"; tail -n 3 gpurun_out/n_streamk.log
  timeout 120 python bench.py --steps 100 --warmup 10 > gpurun_out/n_bench_default.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/n_bench_default.log | cut -c1-200
  DL_B200_TC_STREAMK=1 timeout 120 python bench.py --steps 100 --warmup 10 > gpurun_out/n_bench_streamk.log 2>&1; echo "bench streamk rc=$?"; tail -n 1 gpurun_out/n_bench_streamk.log | cut -c1-200
  timeout 60 python -m torch_distlearn_b200.doctor --selftest > gpurun_out/n_doctor.log 2>&1; echo "doctor rc=$?"; tail -n 6 gpurun_out/n_doctor.log
  timeout 120 python examples/cifar10.py --cuda --numEpochs 1 --batchSize 128 --trainSamples 8192 --testSamples 1024 --graph > gpurun_out/n_example.log 2>&1
  echo "example (native loader on a GPU) rc=$?"; tail -n 4 gpurun_out/n_example.log
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
  timeout 200 $TR --master-port 29533 tests/run_multigpu.py --quick > gpurun_out/n_mg$N.log 2>&1; echo "multigpu rc=$?"; tail -n 5 gpurun_out/n_mg$N.log
  # broadcast with slices owned by the non-root ranks only (was 433 GB/s at 1 GiB on 8 GPUs, NCCL 647)
  timeout 200 $TR --master-port 29541 bench/sweep.py --ops broadcast --min-kb 1024 --max-mb 1024 --out gpurun_out/n_sweep_bc$N.json > gpurun_out/n_sweep_bc$N.log 2>&1
  echo "broadcast sweep rc=$?"; tail -n 12 gpurun_out/n_sweep_bc$N.log | cut -c1-200
  for order in module execution; do
    timeout 120 $TR --master-port 29551 bench.py --gpus $N --steps 100 --warmup 10 --bucket-mb 4 --flat-order $order > gpurun_out/n_bench${N}_b4_$order.log 2>&1
    echo "bench buckets/$order rc=$?"; tail -n 1 gpurun_out/n_bench${N}_b4_$order.log | cut -c1-200
  done
  timeout 120 $TR --master-port 29552 bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/n_bench${N}_b0.log 2>&1; echo "bench rc=$?"; tail -n 1 gpurun_out/n_bench${N}_b0.log | cut -c1-200
fi

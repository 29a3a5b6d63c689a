mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29533 tests/run_multigpu.py --quick > gpurun_out/mg8b.log 2>&1; echo "mg8 rc=$?"; tail -n 2 gpurun_out/mg8b.log
timeout 90 $TR --master-port 29543 bench.py --gpus 8 --steps 100 --warmup 10 > gpurun_out/bench8_tc_b0.log 2>&1; echo "ours b0 rc=$?"; tail -n 1 gpurun_out/bench8_tc_b0.log | cut -c1-200
timeout 90 $TR --master-port 29544 bench.py --gpus 8 --steps 100 --warmup 10 --bucket-mb 4 > gpurun_out/bench8_tc_b4.log 2>&1; echo "ours b4 rc=$?"; tail -n 1 gpurun_out/bench8_tc_b4.log | cut -c1-200
timeout 90 $TR --master-port 29545 bench.py --gpus 8 --steps 100 --warmup 10 --impl nccl > gpurun_out/bench8_nccl2.log 2>&1; echo "nccl rc=$?"; tail -n 1 gpurun_out/bench8_nccl2.log | cut -c1-200
for impl in ours nccl; do
timeout 90 $TR --master-port 29546 bench/bench_ea.py --impl $impl --batch 1 --steps 1000 > gpurun_out/ea8_$impl.log 2>&1; echo "ea $impl rc=$?"; tail -n 1 gpurun_out/ea8_$impl.log | cut -c1-300
done
for impl in ours nccl; do
timeout 90 $TR --master-port 29548 bench/bench_ea.py --impl $impl --batch 64 --tau 1 --steps 500 > gpurun_out/ea8_tau1_$impl.log 2>&1; echo "ea tau1 $impl rc=$?"; tail -n 1 gpurun_out/ea8_tau1_$impl.log | cut -c1-300
done
for impl in ours hostloop; do
timeout 120 $TR --master-port 29547 bench/bench_async.py --impl $impl --steps 100 > gpurun_out/async8_$impl.log 2>&1; echo "async $impl rc=$?"; tail -n 1 gpurun_out/async8_$impl.log | cut -c1-400
done

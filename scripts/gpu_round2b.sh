mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 90 $TR --master-port 29544 bench.py --gpus 2 --steps 50 --warmup 10 --bucket-mb 4 > gpurun_out/bench2_tc_b4.log 2>&1; echo "ours b4 rc=$?"; tail -n 1 gpurun_out/bench2_tc_b4.log | cut -c1-200
timeout 90 $TR --master-port 29543 bench.py --gpus 2 --steps 50 --warmup 10 > gpurun_out/bench2_tc_b0.log 2>&1; echo "ours b0 rc=$?"; tail -n 1 gpurun_out/bench2_tc_b0.log | cut -c1-200
for impl in ours nccl; do
timeout 60 $TR --master-port 29546 bench/bench_ea.py --impl $impl --batch 1 --steps 300 > gpurun_out/ea2_$impl.log 2>&1; echo "ea $impl rc=$?"; tail -n 2 gpurun_out/ea2_$impl.log | cut -c1-300
done
for impl in ours hostloop; do
timeout 90 $TR --master-port 29547 bench/bench_async.py --impl $impl --steps 40 > gpurun_out/async2_$impl.log 2>&1; echo "async $impl rc=$?"; tail -n 2 gpurun_out/async2_$impl.log | cut -c1-400
done
